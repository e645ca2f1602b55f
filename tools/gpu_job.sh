set -x
python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -4 gpurun_out/pytest_gpu.log
python bench.py --steps 3 --warmup 3 > gpurun_out/bench_r1c.json 2> gpurun_out/bench_r1c.err; echo "bench rc=$?"; cat gpurun_out/bench_r1c.json
python bench.py --steps 2 --warmup 3 --wtype q8_0 --no-cpu-baseline > gpurun_out/bench_r1c_q8.json 2> gpurun_out/bench_r1c_q8.err; echo "bench q8 rc=$?"; cat gpurun_out/bench_r1c_q8.json
python bench.py --steps 1 --warmup 3 --frames 4096 --no-cpu-baseline > gpurun_out/bench_r1c_long.json 2> gpurun_out/bench_r1c_long.err; echo "bench long rc=$?"; cat gpurun_out/bench_r1c_long.json
python bench.py --steps 1 --warmup 1 --frames 24 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 200 -c 400 --csv --log-file gpurun_out/launches_r1c.csv python bench.py --steps 1 --warmup 1 --frames 24 --no-cpu-baseline > gpurun_out/ncu1.log 2>&1
