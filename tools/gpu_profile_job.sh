# Round-end measurement job (run with: gpurun --timeout 2400 -- "bash tools/gpu_profile_job.sh"): GPU tests, the bench lines
# of configs 2/3/5, the tcgen05 tap-GEMM check, the ncu launch list and the full captures that profiles/ summarises.
set -x
python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -4 gpurun_out/pytest_gpu.log
python bench.py --steps 3 --warmup 3 > gpurun_out/bench_r1g.json 2> gpurun_out/bench_r1g.err; echo "bench rc=$?"; cat gpurun_out/bench_r1g.json
python bench.py --steps 2 --warmup 3 --wtype q8_0 --no-cpu-baseline > gpurun_out/bench_r1g_q8.json 2> gpurun_out/bench_r1g_q8.err; echo "bench q8 rc=$?"
python bench.py --steps 2 --warmup 3 --wtype q8_0 --temperature 0.9 --top-k 50 --seed 7 --no-cpu-baseline > gpurun_out/bench_r1g_config3.json 2> gpurun_out/bench_r1g_config3.err; echo "bench config3 rc=$?"
python bench.py --steps 1 --warmup 3 --frames 4096 --no-cpu-baseline > gpurun_out/bench_r1g_long.json 2> gpurun_out/bench_r1g_long.err; echo "bench long rc=$?"
timeout 120 tools/microbench/bin/umma_test > gpurun_out/umma_test.txt 2>&1; echo "umma rc=$?"
python bench.py --steps 1 --warmup 1 --frames 24 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches_r1g.csv python bench.py --steps 1 --warmup 1 --frames 24 --no-cpu-baseline > gpurun_out/ncu1.log 2>&1
mkdir -p /tmp/ncu
timeout 600 ncu --set full --clock-control none --import-source on -k regex:lm_kernel -s 40 -c 1 -o /tmp/ncu/lm_kernel python bench.py --steps 1 --warmup 1 --frames 24 --no-cpu-baseline > gpurun_out/ncu2.log 2>&1
ncu -i /tmp/ncu/lm_kernel.ncu-rep --page details > gpurun_out/lm_kernel_r1g_details.txt 2>&1
ncu -i /tmp/ncu/lm_kernel.ncu-rep --page raw --csv > gpurun_out/lm_kernel_r1g_raw.csv 2>&1
timeout 600 ncu --set full --clock-control none -k regex:vgemm_umma -s 130 -c 12 -o /tmp/ncu/vgemm python tools/dev/dev_voc_only.py > gpurun_out/ncu3.log 2>&1
ncu -i /tmp/ncu/vgemm.ncu-rep --page raw --csv > gpurun_out/vgemm_umma_r1g_raw.csv 2>&1
ls -la /tmp/ncu
