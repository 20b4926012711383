#!/bin/bash
# One 8-GPU session: the four BASELINE.json multi-GPU configs + the torch DDP stand-in.  Tight timeouts everywhere.
mkdir -p gpurun_out
export DTC_DIST_TIMEOUT=45 DTC_HANG_DUMP=70
N=${1:-8}
run() { name=$1; shift; timeout 95 "$@" 2> gpurun_out/${name}.err | tail -1 > gpurun_out/${name}.json; echo "$name rc=$? $(cut -c1-260 gpurun_out/${name}.json)"; }
run bench_dp_${N}gpu python bench.py --gpus $N --parallel dp --steps 20 --warmup 3
run bench_tp_${N}gpu python bench.py --gpus $N --parallel tp --steps 20 --warmup 3
run bench_pp_${N}gpu python bench.py --gpus $N --parallel pp --steps 20 --warmup 3
run bench_dptp_medium_${N}gpu python bench.py --gpus $N --parallel dp_tp --steps 10 --warmup 3 --model-config configs/model_config_medium.yaml
run baseline_torch_${N}gpu python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29577 baseline/torch_baseline.py --steps 20
for f in gpurun_out/*_${N}gpu.err; do echo "== $f"; grep -v "^\s*$\|OMP_NUM_THREADS\|\*\*\*\*" $f | tail -4; done
