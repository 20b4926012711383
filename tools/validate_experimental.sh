#!/bin/bash
# First GPU session of the next round: validate and A/B everything that was written after the round-1 GPU budget ran out.
#   gpurun --timeout 420 -- 'bash tools/validate_experimental.sh 1'            (1 GPU: grouped wgrad tests + A/B)
#   gpurun --gpus 2 --timeout 420 -- 'bash tools/validate_experimental.sh 2'   (2 GPUs: PP graphs, TP sequence-parallel tests + A/B)
#   gpurun --gpus 8 --timeout 420 -- 'bash tools/validate_experimental.sh 8'   (8 GPUs: TP8 sequence-parallel vs NCCL, hybrid mesh)
# Every command has its own timeout; one JSON line per bench in gpurun_out/x_*.json, a summary table at the end.
mkdir -p gpurun_out
export DTC_DIST_TIMEOUT=45 DTC_HANG_DUMP=70
N=${1:-1}
run() { name=$1; shift; timeout 100 "$@" 2> gpurun_out/x_${name}.err | tail -1 > gpurun_out/x_${name}.json; echo "$name rc=$?"; }
if [ "$N" = "1" ]; then
  DTC_EXPERIMENTAL=1 timeout 200 python -m pytest tests/test_experimental_gpu.py -q -x 2>&1 | tail -15 | tee gpurun_out/x_experimental_tests.log
  run dp1_default python bench.py --steps 30 --warmup 5
  DTC_WGRAD_GROUPED=1 run dp1_wgrad_grouped python bench.py --steps 30 --warmup 5
  # the fairer bar for the kernels: the same cuBLAS / SDPA / fused-AdamW step without host launch gaps
  timeout 100 python baseline/torch_baseline.py --steps 30 --graph 2> gpurun_out/x_torch_graph.err | tail -1 > gpurun_out/baseline_torch_graph_1gpu.json
  echo "torch graph baseline rc=$? $(cut -c1-300 gpurun_out/baseline_torch_graph_1gpu.json)"
fi
if [ "$N" = "2" ]; then
  DTC_EXPERIMENTAL=1 timeout 300 python -m pytest tests/test_multigpu.py -q -x -k "sequence_parallel or cuda_graph_per_action or tp2 or folded" 2>&1 | tail -15 | tee gpurun_out/x_multigpu_tests.log
  run pp2_eager python bench.py --gpus 2 --parallel pp --steps 30 --warmup 5
  run pp2_graph python bench.py --gpus 2 --parallel pp --steps 30 --warmup 5 --pp-graph
  run tp2_seqpar python bench.py --gpus 2 --parallel tp --steps 30 --warmup 5
  run tp2_nccl python bench.py --gpus 2 --parallel tp --steps 30 --warmup 5 --no-tp-seq-parallel
  DTC_SP_FOLD=1 run tp2_seqpar_fold python bench.py --gpus 2 --parallel tp --steps 30 --warmup 5
  DTC_WGRAD_GROUPED=1 run dp2_wgrad_grouped python bench.py --gpus 2 --steps 30 --warmup 5
fi
if [ "$N" -ge 4 ]; then
  run tp${N}_seqpar python bench.py --gpus $N --parallel tp --steps 30 --warmup 5
  run tp${N}_nccl python bench.py --gpus $N --parallel tp --steps 30 --warmup 5 --no-tp-seq-parallel
  DTC_SP_FOLD=1 run tp${N}_seqpar_fold python bench.py --gpus $N --parallel tp --steps 30 --warmup 5
  run pp${N}_eager python bench.py --gpus $N --parallel pp --steps 30 --warmup 5
  run pp${N}_graph python bench.py --gpus $N --parallel pp --steps 30 --warmup 5 --pp-graph
  if [ "$N" = "8" ]; then
    DTC_SP_HYBRID=1 run dptp_medium_seqpar python bench.py --gpus 8 --parallel dp_tp --steps 10 --warmup 3 --model-config configs/model_config_medium.yaml
    run dptp_medium_nccl python bench.py --gpus 8 --parallel dp_tp --steps 10 --warmup 3 --model-config configs/model_config_medium.yaml --no-tp-seq-parallel
  fi
fi
if [ "$N" -ge 2 ]; then
  timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29581 \
    tools/comm_bench.py 2> gpurun_out/x_comm_bench_${N}gpu.err | tail -1 > gpurun_out/comm_bench_${N}gpu_r2.json
  echo "comm_bench rc=$? $(cut -c1-600 gpurun_out/comm_bench_${N}gpu_r2.json)"
fi
python - <<'PY'
import glob, json
for f in sorted(glob.glob("gpurun_out/x_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f"{f.split('/')[-1]:36s} {d['config']['parallelism']:8s} {d['ms_per_step']:8.3f} ms  {d['value']:12.0f} tok/s  "
              f"loss {d.get('loss_first', 0):.4f} -> {d.get('loss_last', 0):.4f}  tp={d['config']['comm_path']['tp'][:40]}")
    except Exception as e:
        print(f"{f}: no result ({type(e).__name__}); see {f[:-5]}.err")
PY
