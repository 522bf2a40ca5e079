#!/bin/bash
# round 2, GPU call 1: baseline tests + bench, element-wise bisect of the GEMM epilogue item over build variants, racecheck
mkdir -p gpurun_out
V=embeddings.cpp_b200/lib/variants
nvidia-smi -L
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2c1_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2c1_pytest.log
: > gpurun_out/r2c1_race.log
for v in default ldx1 ldx1_xcta ldx1_waitall ldx1_delay ldx1_sa4sb4 ldx1_sink xcta; do
  if [ $v = default ]; then lib=$PWD/embeddings.cpp_b200/lib/libbert.so; else lib=$PWD/$V/libbert_$v.so; fi
  echo "=== variant $v" >> gpurun_out/r2c1_race.log
  BERT_B200_LIB=$lib timeout 300 python tools/gemm_race.py 65536,2304,768,2,1,2 65536,3072,768,2,43,2 >> gpurun_out/r2c1_race.log 2>&1
  echo "race $v rc=$?"
done
: > gpurun_out/r2c1_e2e.log
for v in ldx1 ldx1_xcta ldx1_waitall ldx1_sa4sb4; do
  echo "=== variant $v" >> gpurun_out/r2c1_e2e.log
  BERT_B200_LIB=$PWD/$V/libbert_$v.so timeout 300 python -m pytest tests/test_gpu.py -m gpu -x -q -k full_size 2>&1 | grep -E "^E  .*assert|passed|failed" | head -4 >> gpurun_out/r2c1_e2e.log
done
timeout 600 python bench.py > gpurun_out/r2c1_bench.json 2> gpurun_out/r2c1_bench.err; echo "bench rc=$?"
BERT_GEMM_MAX_CLUSTERS=2 timeout 700 compute-sanitizer --tool racecheck --print-limit 40 python tools/sanitize_cases.py gemm attn > gpurun_out/r2c1_racecheck.log 2>&1; echo "racecheck rc=$?"
tail -5 gpurun_out/r2c1_racecheck.log
