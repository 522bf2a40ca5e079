#!/bin/bash
# round 2 GPU call: the edge-case test with the single-string form of BertModel.encode
mkdir -p gpurun_out
timeout 60 python -m pytest tests/test_gpu.py -m gpu -x -q -k "edge_cases" 2>&1 | tail -2 | cut -c1-300
