# N-GPU validation (N = number of visible GPUs): the multi-GPU tests that need > 2 GPUs, every bench arm, the torch bar
mkdir -p gpurun_out
N=$(python -c "import torch;print(torch.cuda.device_count())")
export DTC_DIST_TIMEOUT=90 DTC_EXPERIMENTAL=1
if [ "$N" = "4" ]; then K="tp4 or dp2_tp2"; else K="tp8 or dp8 or pp8"; fi
timeout 600 python -m pytest tests/test_multigpu.py -v -x -k "$K" 2>&1 | grep -E "PASSED|FAILED|SKIPPED|ERROR|passed|failed|Error" | tail -20 | tee gpurun_out/r2_pytest_multigpu_${N}gpu.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 30 --warmup 5 --arm-timeout 100 > gpurun_out/r2_bench$N.json 2> gpurun_out/r2_bench$N.err; echo bench rc=$?
python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/r2_bench$N.json") if l.startswith("{")][-1])
print("dp", round(d["ms_per_step"],3), d["loss_last"])
for k,v in d.get("extra",{}).items(): print(k, round(v.get("ms_per_step",0),3) if "ms_per_step" in v else v)
PY
for mode in weak strong; do
  if [ $mode = weak ]; then B=8; else B=$((8/N)); fi
  timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29551 baseline/torch_baseline.py --steps 20 --graph --batch $B 2> gpurun_out/r2_torch_${mode}_$N.err | tail -1 > gpurun_out/baseline_torch_graph_${mode}_${N}gpu.json; echo "torch $mode rc=$? $(cut -c1-220 gpurun_out/baseline_torch_graph_${mode}_${N}gpu.json)"
done
if [ "$N" = "8" ]; then
  DTC_SP_HYBRID=1 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus 8 --parallel dp_tp --steps 10 --warmup 3 --model-config configs/model_config_medium.yaml > gpurun_out/r2_dptp_medium_seqpar.json 2> gpurun_out/r2_dptp_medium_seqpar.err; echo "hybrid seqpar rc=$? $(cut -c1-200 gpurun_out/r2_dptp_medium_seqpar.json)"
fi
