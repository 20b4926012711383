mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
b() { name=$1; shift; env "$@" timeout 120 python bench.py --steps 30 --warmup 5 > gpurun_out/r2_b1_$name.json 2>gpurun_out/r2_b1_$name.err; python -c "import json;d=json.load(open('gpurun_out/r2_b1_$name.json'));print('$name',round(d['ms_per_step'],3),round(d['e2e']['ms_per_step'],3),d['loss_last'])"; }
b default A=1
