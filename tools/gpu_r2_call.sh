#!/bin/bash
# round 2 GPU call: the per-kernel attention check against numpy
mkdir -p gpurun_out
timeout 60 python -m pytest tests/test_gpu.py -m gpu -x -q -k "attention_against_numpy" 2>&1 | tail -4 | cut -c1-400
