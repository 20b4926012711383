mkdir -p gpurun_out
export DTC_DIST_TIMEOUT=90
r() { name=$1; shift; timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 30 --warmup 5 "$@" > gpurun_out/r2_$name.json 2> gpurun_out/r2_$name.err; python -c "import json;d=json.loads([l for l in open('gpurun_out/r2_$name.json') if l.startswith('{')][-1]);print('$name',round(d['ms_per_step'],3),round(d['e2e']['ms_per_step'],3),d['loss_last'],d['config'].get('pp_layers_per_stage'))"; }
r pp2_even --parallel pp
r pp2_balanced --parallel pp --pp-balanced
r pp2_balanced_m4_1f1b --parallel pp --pp-balanced --pp-microbatches 4 --pp-schedule 1f1b
for mode in weak strong; do
  if [ $mode = weak ]; then B=8; else B=4; fi
  timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29551 baseline/torch_baseline.py --steps 30 --graph --batch $B 2> gpurun_out/r2_torch_${mode}_2.err | tail -1 > gpurun_out/baseline_torch_graph_${mode}_2gpu.json; echo "torch $mode $(cut -c100-330 gpurun_out/baseline_torch_graph_${mode}_2gpu.json)"
done
