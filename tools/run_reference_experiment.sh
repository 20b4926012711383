#!/bin/bash
# The reference's own experiment (README: run each strategy for 5000 steps on identical data, compare loss and time)
# on N GPUs of one box: outputs/{dp,tp,pp}/log.csv + outputs/summary.csv, as in the reference's committed outputs/.
N=${1:-2}
STEPS=${2:-5000}
export DTC_DIST_TIMEOUT=90
mkdir -p gpurun_out/outputs
for s in dp tp pp; do
  timeout 300 python main.py --train_config_path configs/train_config_$s.yaml --num_devices $N --set steps=$STEPS \
      --set output_dir=gpurun_out/outputs/$s 2>gpurun_out/outputs/$s.err | tail -4
done
# PP again with the same initial weights as DP/TP and a global clip (parity variant)
timeout 300 python main.py --train_config_path configs/train_config_pp.yaml --num_devices $N --set steps=$STEPS \
    --set output_dir=gpurun_out/outputs/pp_shared --set pp_init=shared --set pp_clip=global 2>gpurun_out/outputs/pp_shared.err | tail -2
python plot.py gpurun_out/outputs/
