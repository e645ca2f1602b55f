set -x
python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -4 gpurun_out/pytest_gpu.log
python bench.py --steps 3 --warmup 3 > gpurun_out/bench_r1f.json 2> gpurun_out/bench_r1f.err; echo "bench rc=$?"; cat gpurun_out/bench_r1f.json
python bench.py --steps 2 --warmup 3 --wtype q8_0 --no-cpu-baseline > gpurun_out/bench_r1f_q8.json 2> gpurun_out/bench_r1f_q8.err; echo "bench q8 rc=$?"
python bench.py --steps 1 --warmup 3 --frames 4096 --no-cpu-baseline > gpurun_out/bench_r1f_long.json 2> gpurun_out/bench_r1f_long.err; echo "bench long rc=$?"
python bench.py --steps 1 --warmup 1 --frames 24 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches_r1f.csv python bench.py --steps 1 --warmup 1 --frames 24 --no-cpu-baseline > gpurun_out/ncu1.log 2>&1
mkdir -p /tmp/ncu
timeout 600 ncu --set full --clock-control none --import-source on -k regex:lm_kernel -s 40 -c 1 -o /tmp/ncu/lm_kernel python bench.py --steps 1 --warmup 1 --frames 24 --no-cpu-baseline > gpurun_out/ncu2.log 2>&1
ncu -i /tmp/ncu/lm_kernel.ncu-rep --page details > gpurun_out/lm_kernel_r1f_details.txt 2>&1
ncu -i /tmp/ncu/lm_kernel.ncu-rep --page raw --csv > gpurun_out/lm_kernel_r1f_raw.csv 2>&1
ls -la /tmp/ncu
