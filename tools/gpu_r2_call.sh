#!/bin/bash
# round 2 GPU call: last sanity run of the shipped library, part 2 (all GEMM self-tests, full-size properties and determinism)
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu.py -m gpu -x -q -k "gemm or full_size" > gpurun_out/r2_sanity2_pytest.log 2>&1; tail -2 gpurun_out/r2_sanity2_pytest.log | cut -c1-300
