timeout 300 python -m pytest tests/test_kernels_gpu.py tests/test_engine_gpu.py -q -x 2>&1 | tail -4
DTC_ATTN_BWD_PERSIST=1 timeout 100 python -m pytest tests/test_kernels_gpu.py tests/test_zz_no_alloc.py -q -x -k "attention or static" 2>&1 | tail -3
b() { name=$1; shift; env "$@" timeout 120 python bench.py --steps 30 --warmup 5 > gpurun_out/r2_b1_$name.json 2>gpurun_out/r2_b1_$name.err; python -c "import json;d=json.load(open('gpurun_out/r2_b1_$name.json'));print('$name',round(d['ms_per_step'],3),round(d['e2e']['ms_per_step'],3),d['loss_last'])"; }
b default A=1
b nopairlayers DTC_GEMM_PAIR_LAYERS=0
b bwd1cta DTC_BWD_GEMM_IMPL=4
b persist DTC_ATTN_BWD_PERSIST=1
timeout 100 python tools/attn_bench.py 2>&1 | tee gpurun_out/r2_attn_bench_pipe.txt
DTC_ATTN_BWD_PERSIST=1 timeout 100 python tools/attn_bench.py 2>&1 | tee gpurun_out/r2_attn_bench_persist.txt
export DTC_ATTN_BWD_PERSIST=1
bash tools/profile_r2.sh 2>&1 | grep -E "Report|error|rc=" 
