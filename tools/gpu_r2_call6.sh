#!/bin/bash
# round 2, GPU call 6: the fixed kernel (4-bit slot handed back at the end of the stage; real load fence before the residual
# buffer goes back to the TMA; both epilogue TMEM loads up front) -- every instantiation, perturbed builds, the full suite, bench
mkdir -p gpurun_out
V=embeddings.cpp_b200/lib/variants
timeout 900 python tools/gemm_race.py --matrix --brief > gpurun_out/r2c6_matrix.log 2>&1; echo "matrix rc=$?"
grep -c "off vs validation kernel 0 of" gpurun_out/r2c6_matrix.log; grep "off vs" gpurun_out/r2c6_matrix.log | grep -v "kernel 0 of" | head
: > gpurun_out/r2c6_race.log
for v in ldx0 dq1 sa4sb4 race_r1; do
  echo "=== variant $v" >> gpurun_out/r2c6_race.log
  BERT_B200_LIB=$PWD/$V/libbert_$v.so timeout 300 python tools/gemm_race.py --brief 65536,2304,768,2,1,6 65536,768,768,2,29,3 >> gpurun_out/r2c6_race.log 2>&1
  echo "race $v rc=$?"
done
grep -E "variant|case|run " gpurun_out/r2c6_race.log | cut -c1-170
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2c6_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2c6_pytest.log
: > gpurun_out/r2c6_e2e.log
for v in ldx0 dq1 dq1 sa4sb4 sa4sb4 race_r1; do
  echo "=== variant $v" >> gpurun_out/r2c6_e2e.log
  BERT_B200_LIB=$PWD/$V/libbert_$v.so timeout 300 python -m pytest tests/test_gpu.py -m gpu -x -q -k full_size 2>&1 | grep -E "^E  .*assert|passed|failed" | head -4 >> gpurun_out/r2c6_e2e.log
done
cat gpurun_out/r2c6_e2e.log
for r in 1 2; do
  timeout 600 python bench.py --no-cpu-baseline 2>/dev/null | grep '^{' > gpurun_out/r2c6_bench_default_$r.json
  BERT_B200_LIB=$PWD/$V/libbert_ldx0.so timeout 600 python bench.py --no-cpu-baseline 2>/dev/null | grep '^{' > gpurun_out/r2c6_bench_ldx0_$r.json
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2c6_bench_*.json')):
    d=json.load(open(f)); print(f, round(d['value'],1), round(d['e2e']['value'],1), round(d['roofline']['achieved'],1), d['roofline']['by_kernel_ms'], d['clocks']['sm_mhz'])
PY
