#!/bin/bash
# ONE ncu capture per gpurun call (1 GPU).  Usage:  gpurun --timeout 600 -- 'bash tools/profile_round2.sh <what>'
#   launches   every kernel of ~2 training steps with its device time        -> gpurun_out/launches_r2.csv
#   attn_bwd   ncu --set full on the attention backward kernel               -> gpurun_out/prof_attn_bwd.ncu-rep
#   attn_fwd   ncu --set full on the attention forward kernel                -> gpurun_out/prof_attn_fwd.ncu-rep
#   gemm       ncu --set full on the persistent GEMM (qkv / fc1 / fc2 shapes) -> gpurun_out/prof_gemm.ncu-rep
#   wgrad      ncu --set full on the grouped wgrad kernel                     -> gpurun_out/prof_wgrad.ncu-rep
# Read here with:  ncu -i gpurun_out/<name>.ncu-rep --page raw --csv | grep -E 'gpu__time_duration.sum|sm__pipe_tensor|lts__t_bytes|dram__bytes|sm__warps_active|launch__registers'
#                  ncu -i gpurun_out/<name>.ncu-rep --page source --csv      (stall reasons per line; the library is built with -lineinfo)
#                  python tools/ncu_launch_table.py gpurun_out/launches_r2.csv --steps 2.1 --md profiles/r2/step_budget_ncu.md   (steps = launches / ~300)
mkdir -p gpurun_out
what=${1:-launches}
case "$what" in
  launches)
    cmd="python bench.py --steps 2 --warmup 3 --no-graph"
    $cmd > gpurun_out/plain.log 2>&1 || { tail -20 gpurun_out/plain.log; exit 1; }
    # ~300 launches per step without the graph: skip the 3 warm-up steps, keep about two steps
    ncu --metrics gpu__time_duration.sum --clock-control none -s 900 -c 650 --csv --log-file gpurun_out/launches_r2.csv $cmd \
      > gpurun_out/ncu.log 2>&1; echo "ncu rc=$?"; grep -c gpu__time_duration gpurun_out/launches_r2.csv ;;
  attn_bwd|attn_fwd|gemm)
    case "$what" in attn_bwd) pat="flash_attn_bwd"; tgt=attn;; attn_fwd) pat="flash_attn_fwd"; tgt=attn;; gemm) pat="gemm_persistent"; tgt=gemm;; esac
    cmd="python tools/profile_target.py $tgt"
    $cmd > gpurun_out/plain.log 2>&1 || { tail -20 gpurun_out/plain.log; exit 1; }
    ncu --set full --clock-control none --import-source on -k regex:$pat -s 3 -c 3 -o gpurun_out/prof_$what -f $cmd \
      > gpurun_out/ncu.log 2>&1; echo "ncu rc=$?"; tail -5 gpurun_out/ncu.log ;;
  wgrad)
    cmd="env DTC_EXPERIMENTAL=1 python -m pytest tests/test_experimental_gpu.py -q -x -k wgrad_grouped"
    $cmd > gpurun_out/plain.log 2>&1 || { tail -20 gpurun_out/plain.log; exit 1; }
    ncu --set full --clock-control none --import-source on -k regex:gemm_grouped_wgrad -c 2 -o gpurun_out/prof_wgrad -f $cmd \
      > gpurun_out/ncu.log 2>&1; echo "ncu rc=$?"; tail -5 gpurun_out/ncu.log ;;
  *) echo "unknown target $what"; exit 2 ;;
esac
ls -la gpurun_out/*.ncu-rep gpurun_out/launches_r2.csv 2>/dev/null
