#!/bin/bash
# round 2 GPU call (2 GPUs): the in-process sharding test and the torchrun bench line on the final tree
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu.py -m gpu -x -q -k "in_process_sharding" 2>&1 | tail -2 | cut -c1-200
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | grep '^{' > gpurun_out/bench_r2_last_2gpu.json
python - <<PY
import json
d=json.load(open('gpurun_out/bench_r2_last_2gpu.json')); print('n=2', 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), {k:v for k,v in d.items() if 'inprocess' in k or k=='n_gpus'})
PY
