# final 1-GPU validation of the committed build: tests, default bench, the other BASELINE configs
(timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3) 
timeout 300 python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err; tail -c 2500 gpurun_out/bench_final.json
for w in bge-small-f16-b256-s128 bge-large-q4_0-b1024-s256 minilm-q4_0-b1738-var; do
  timeout 300 python bench.py --workload $w --steps 10 --warmup 3 > gpurun_out/bench_final_$w.json 2>/dev/null; python - <<PY
import json
d=json.loads(open("gpurun_out/bench_final_$w.json").read().strip().splitlines()[-1])
print("$w", round(d["value"],1), "e2e", round(d["e2e"]["value"],1), "gemm TF", round(d["roofline"]["achieved"],1), d["cpu_baseline"])
PY
done
