#!/bin/bash
# round 2 GPU call: golden parity through the C ABI with the final host tokenizer
timeout 25 python -m pytest tests/test_gpu.py -m gpu -x -q -k "golden" 2>&1 | tail -2 | cut -c1-300
