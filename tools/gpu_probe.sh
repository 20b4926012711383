#!/bin/bash
# Runs the GPU kernel tests group by group, each in its own process (a trapped kernel poisons the CUDA
# context of its process only) and under `timeout`.  Logs land in gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
status=0
for grp in "library" "gemm_majors" "gemm_large or gemm_epilogues or gemm_wgrad or gemm_strided" "attention" "layernorm" "embed" "cross_entropy" "adamw or colsum"; do
  tag=$(echo "$grp" | tr ' ' '_' | cut -c1-24)
  echo "=== group: $grp"
  timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -q -k "$grp" -p no:cacheprovider > gpurun_out/probe_$tag.log 2>&1
  rc=$?
  tail -n 15 gpurun_out/probe_$tag.log
  echo "rc=$rc"
  [ $rc -ne 0 ] && status=1
done
exit $status
