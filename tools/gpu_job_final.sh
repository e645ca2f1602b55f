set -x
python bench.py > gpurun_out/bench_r1g.json 2> gpurun_out/bench_r1g.err; echo "bench rc=$?"
python bench.py --steps 2 --warmup 3 --wtype q8_0 --no-cpu-baseline > gpurun_out/bench_r1g_q8.json 2> gpurun_out/bench_r1g_q8.err; echo "bench q8 rc=$?"
python bench.py --steps 2 --warmup 3 --wtype q8_0 --temperature 0.9 --top-k 50 --seed 7 --no-cpu-baseline > gpurun_out/bench_r1g_config3.json 2> gpurun_out/bench_r1g_config3.err; echo "bench config3 rc=$?"
python bench.py --steps 1 --warmup 3 --frames 4096 --no-cpu-baseline > gpurun_out/bench_r1g_long.json 2> gpurun_out/bench_r1g_long.err; echo "bench long rc=$?"
timeout 300 python tests/dev_phase_profile.py 40 > gpurun_out/phase_b1.txt 2>&1
timeout 300 python tests/dev_phase_profile.py 40 8 > gpurun_out/phase_b8.txt 2>&1
for f in gpurun_out/bench_r1g*.json; do python -c "
import json,sys; d=json.load(open('$f')); print('$f', round(d['value'],2), round(d['roofline']['frac'],4), d['device_ms_per_step'], d.get('batch64',{}).get('value'))"; done
