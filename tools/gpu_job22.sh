set -x
timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -q -x -k "sampler or error_behaviour or greedy_codes" 2>&1 | tail -3
python bench.py --steps 2 --warmup 3 --wtype q8_0 --temperature 0.9 --top-k 50 --seed 7 --no-cpu-baseline > gpurun_out/bench_r1g_config3.json 2> gpurun_out/bench_r1g_config3.err; echo "bench config3 rc=$?"; cat gpurun_out/bench_r1g_config3.json | cut -c1-200; python -c "
import json; d=json.load(open('gpurun_out/bench_r1g_config3.json')); print(d['value'], d['device_ms_per_step'])"
