#!/bin/bash
# N-GPU validation in ONE gpurun call (N = number of visible GPUs: 2, 4 or 8; charged N x the box time, so keep it short):
#   gpurun --gpus N --timeout 1500 -- 'bash tools/validate_multigpu.sh'
# runs the multi-GPU tests that need exactly this box size, every bench arm (DP global batch 8, DP weak, TP = N, PP = N),
# the graphed torch DDP bar (weak: batch 8 per rank; strong: global batch 8) and, on 8 GPUs, the hybrid DP2 x TP4 medium model.
# Logs: gpurun_out/pytest_multigpu_${N}gpu.log, bench_all_arms_${N}gpu.json, baseline_torch_graph_{weak,strong}_${N}gpu.json.
mkdir -p gpurun_out
N=$(python -c "import torch;print(torch.cuda.device_count())")
export DTC_DIST_TIMEOUT=90
case "$N" in 2) K="not (tp4 or tp8 or dp8 or pp8 or dp2_tp2)";; 4) K="tp4 or dp2_tp2";; *) K="tp8 or dp8 or pp8";; esac
timeout 900 python -m pytest tests/test_multigpu.py -v -x -k "$K" 2>&1 | grep -E "PASSED|FAILED|SKIPPED|ERROR|passed|failed|Error" | tail -40 | tee gpurun_out/pytest_multigpu_${N}gpu.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 30 --warmup 5 --arm-timeout 100 > gpurun_out/bench_all_arms_${N}gpu.json 2> gpurun_out/bench_all_arms_${N}gpu.err; echo "bench rc=$?"
python - <<PY
import json
d = json.loads([l for l in open("gpurun_out/bench_all_arms_${N}gpu.json") if l.startswith("{")][-1])
print("dp", round(d["ms_per_step"], 3), d["loss_last"])
for k, v in d.get("extra", {}).items():
    print(k, round(v["ms_per_step"], 3) if "ms_per_step" in v else v)
PY
for mode in weak strong; do
  if [ $mode = weak ]; then B=8; else B=$((8 / N)); fi
  timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29551 baseline/torch_baseline.py --steps 20 --graph --batch $B 2> gpurun_out/torch_${mode}_$N.err | tail -1 > gpurun_out/baseline_torch_graph_${mode}_${N}gpu.json
  echo "torch $mode rc=$? $(cut -c1-220 gpurun_out/baseline_torch_graph_${mode}_${N}gpu.json)"
done
if [ "$N" = "8" ]; then
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus 8 --parallel dp_tp --steps 10 --warmup 3 --model-config configs/model_config_medium.yaml > gpurun_out/bench_dptp_medium_8gpu.json 2> gpurun_out/bench_dptp_medium_8gpu.err
  echo "hybrid DP2xTP4 medium rc=$? $(cut -c1-200 gpurun_out/bench_dptp_medium_8gpu.json)"
fi
