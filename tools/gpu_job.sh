set -x
python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -4 gpurun_out/pytest_gpu.log
python bench.py --steps 3 --warmup 3 > gpurun_out/bench_r1e.json 2> gpurun_out/bench_r1e.err; echo "bench rc=$?"; cat gpurun_out/bench_r1e.json
python bench.py --steps 1 --warmup 1 --frames 24 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches_r1e.csv python bench.py --steps 1 --warmup 1 --frames 24 --no-cpu-baseline > gpurun_out/ncu1.log 2>&1
tail -2 gpurun_out/ncu1.log
