#!/bin/bash
# ncu --set full captures of the hot kernels (1 GPU, one gpurun call).  Each target is first run plainly (must exit 0).
#   gpurun --timeout 900 -- 'bash tools/profile_r2.sh'
# Read here:  ncu -i gpurun_out/prof_<name>.ncu-rep --page raw --csv | python tools/ncu_summary.py
mkdir -p gpurun_out
set -x
python tools/profile_target.py attn > gpurun_out/plain_attn.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:flash_attn_fwd -s 2 -c 1 -o gpurun_out/prof_attn_fwd -f python tools/profile_target.py attn > gpurun_out/ncu_attn_fwd.log 2>&1
tail -2 gpurun_out/ncu_attn_fwd.log
ncu --set full --clock-control none --import-source on -k regex:flash_attn_bwd -s 2 -c 1 -o gpurun_out/prof_attn_bwd -f python tools/profile_target.py attn > gpurun_out/ncu_attn_bwd.log 2>&1
tail -2 gpurun_out/ncu_attn_bwd.log
DTC_PROFILE_IMPL=3 python tools/profile_target.py gemm > gpurun_out/plain_gemm.log 2>&1 &&
DTC_PROFILE_IMPL=3 ncu --set full --clock-control none --import-source on -k regex:gemm_2cta -s 8 -c 4 -o gpurun_out/prof_gemm_pair -f python tools/profile_target.py gemm > gpurun_out/ncu_gemm_pair.log 2>&1
tail -2 gpurun_out/ncu_gemm_pair.log
ls -la gpurun_out/*.ncu-rep
