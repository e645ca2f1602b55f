set -x
python bench.py > gpurun_out/bench_r1g.json 2> gpurun_out/bench_r1g.err; echo "bench rc=$?"
python bench.py --steps 1 --warmup 3 --frames 4096 --no-cpu-baseline > gpurun_out/bench_r1g_long.json 2> gpurun_out/bench_r1g_long.err; echo "bench long rc=$?"
timeout 300 python tools/dev/dev_phase_profile.py 40 > gpurun_out/phase_b1.txt 2>&1
timeout 600 python -m pytest tests/test_parity_gpu.py tests/test_longform_gpu.py -m gpu -q -x 2>&1 | tail -2
