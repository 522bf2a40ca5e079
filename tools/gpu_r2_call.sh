#!/bin/bash
# round 2 GPU call on 2 GPUs: new per-kernel numpy check, in-process sharding test, the driver's N=2 bench command (both arms)
mkdir -p gpurun_out
echo "== numpy GEMM check + sharding"; timeout 400 python -m pytest tests/test_gpu.py -m gpu -x -q -k "against_numpy or in_process_sharding" 2>&1 | tail -3 | cut -c1-300
echo "== torchrun bench, 2 ranks"; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 10 --warmup 3 2>gpurun_out/r2_n2.err | grep '^{' > gpurun_out/bench_r2_final_2gpu.json
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_r2_final_2gpu.json')); print('n_gpus', d['n_gpus'], 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), d.get('e2e_inprocess'), d['clocks'])
PY
echo "== reference arm, 2 ranks"; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 2>/dev/null | grep '^{' | cut -c1-400
