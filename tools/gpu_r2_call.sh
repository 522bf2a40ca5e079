#!/bin/bash
# round 2 GPU call on 8 GPUs: in-process sharding test, then BASELINE config 5 as written (1 M sentences, strong scaling)
mkdir -p gpurun_out
nvidia-smi -L | head -8
echo "== in-process sharding test"; timeout 300 python -m pytest tests/test_gpu.py -m gpu -x -q -k "in_process_sharding" 2>&1 | tail -3 | cut -c1-300
bash tools/gpu_r2_1m.sh 8
