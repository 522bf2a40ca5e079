#!/bin/bash
# round 2 GPU call on N GPUs: in-process sharding test, then the driver's bench command at N ranks (value, e2e, e2e_inprocess)
N=${1:-8}
mkdir -p gpurun_out
echo "== sharding test"; timeout 300 python -m pytest tests/test_gpu.py -m gpu -x -q -k "in_process_sharding" 2>&1 | tail -2 | cut -c1-300
echo "== torchrun bench, $N ranks"; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 10 --warmup 3 2>gpurun_out/r2_n$N.err | grep '^{' > gpurun_out/bench_r2_final_${N}gpu.json
python - <<PY
import json
d=json.load(open('gpurun_out/bench_r2_final_${N}gpu.json')); print('n_gpus', d['n_gpus'], 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), d.get('e2e_inprocess'), d['clocks'])
PY
