#!/bin/bash
# BASELINE config 5 as written: ONE job of 1 M synthetic sentences x 256 tokens, bge-large q4_0, strong scaling over $1 GPUs
N=${1:-1}
mkdir -p gpurun_out
if [ "$N" = 1 ]; then
  timeout 1500 python bench.py --workload bge-large-q4_0-1M-s256 --no-cpu-baseline 2>gpurun_out/r2_1m_n1.err | grep '^{' > gpurun_out/bench_r2_1M_n1.json
else
  timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --workload bge-large-q4_0-1M-s256 --no-cpu-baseline 2>gpurun_out/r2_1m_n$N.err | grep '^{' > gpurun_out/bench_r2_1M_n$N.json
fi
python - <<PY
import json
d=json.load(open('gpurun_out/bench_r2_1M_n$N.json'))
print('n_gpus', d['n_gpus'], 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'ms', round(d['ms_per_step'],1), d.get('e2e_inprocess'), d['host'], d['clocks'])
PY
tail -3 gpurun_out/r2_1m_n$N.err | cut -c1-300
