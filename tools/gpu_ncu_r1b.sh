# ncu captures of the HEAD build: launch list of the default bench, full-set + source of attention and the layer GEMMs
mkdir -p gpurun_out
timeout 300 python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/bench_pre_ncu.json 2> /dev/null || exit 1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r1_head.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
timeout 300 python tools/prof_attn.py > /dev/null 2>&1 && timeout 600 ncu --set full --import-source on --clock-control none -k regex:attention_tc -c 2 -f -o gpurun_out/prof_r1_head_attn python tools/prof_attn.py > gpurun_out/ncu_attn.log 2>&1
timeout 300 python tools/prof_kernels.py 131072 qkv,o_ln,ffn_in_h2,ffn_out_ln 1 > /dev/null 2>&1 && timeout 900 ncu --set full --import-source on --clock-control none -k regex:gemm_tc_kernel -c 8 -f -o gpurun_out/prof_r1_head_gemms python tools/prof_kernels.py 131072 qkv,o_ln,ffn_in_h2,ffn_out_ln 1 > gpurun_out/ncu_gemms.log 2>&1
ls -la gpurun_out/
