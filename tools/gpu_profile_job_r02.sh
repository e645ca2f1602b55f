# Round-2 measurement job: gpurun --timeout 2400 -- "bash tools/gpu_profile_job_r02.sh"
# bench lines of configs 2/3/5, phase profiles, the ncu launch list and the full captures that profiles/r02_* summarise.
set -x
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_r02.json 2> gpurun_out/bench_r02.err; echo "bench rc=$?"
python bench.py --steps 2 --warmup 3 --wtype q8_0 --no-cpu-baseline --no-batch > gpurun_out/bench_r02_q8.json 2> gpurun_out/bench_r02_q8.err; echo "bench q8 rc=$?"
python bench.py --steps 2 --warmup 3 --wtype q8_0 --temperature 0.9 --top-k 50 --seed 7 --no-cpu-baseline --no-batch > gpurun_out/bench_r02_config3.json 2> gpurun_out/bench_r02_config3.err; echo "bench config3 rc=$?"
python bench.py --steps 1 --warmup 3 --frames 4096 --no-cpu-baseline --no-batch > gpurun_out/bench_r02_long.json 2> gpurun_out/bench_r02_long.err; echo "bench long rc=$?"
timeout 300 python tools/dev/dev_phase_profile.py 40 > gpurun_out/phase_r02_batch1.txt 2>&1
Q3T_WTYPE=q8_0 timeout 300 python tools/dev/dev_phase_profile.py 40 > gpurun_out/phase_r02_q8.txt 2>&1
Q3T_BATCH_WAVE=8 timeout 300 python tools/dev/dev_phase_profile.py 40 8 > gpurun_out/phase_r02_batch8.txt 2>&1
timeout 300 python tools/dev/dev_phase_profile.py 40 16 > gpurun_out/phase_r02_batch16.txt 2>&1
python bench.py --steps 1 --warmup 1 --frames 24 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches_r02.csv python bench.py --steps 1 --warmup 1 --frames 24 --no-cpu-baseline > gpurun_out/ncu1.log 2>&1
mkdir -p /tmp/ncu
timeout 600 ncu --set full --clock-control none --import-source on -k regex:lm_kernel -s 40 -c 1 -o /tmp/ncu/lm_kernel python bench.py --steps 1 --warmup 1 --frames 24 --no-cpu-baseline --no-batch > gpurun_out/ncu2.log 2>&1
ncu -i /tmp/ncu/lm_kernel.ncu-rep --page details > gpurun_out/lm_kernel_r02_details.txt 2>&1
ncu -i /tmp/ncu/lm_kernel.ncu-rep --page raw --csv > gpurun_out/lm_kernel_r02_raw.csv 2>&1
timeout 600 ncu --set full --clock-control none -k regex:lm_kernel -s 40 -c 1 -o /tmp/ncu/lm_kernel_q8 python bench.py --steps 1 --warmup 1 --frames 24 --wtype q8_0 --no-cpu-baseline --no-batch > gpurun_out/ncu2q.log 2>&1
ncu -i /tmp/ncu/lm_kernel_q8.ncu-rep --page raw --csv > gpurun_out/lm_kernel_q8_r02_raw.csv 2>&1
timeout 600 ncu --set full --clock-control none -k regex:vgemm_umma -s 130 -c 12 -o /tmp/ncu/vgemm python tools/dev/dev_voc_only.py > gpurun_out/ncu3.log 2>&1
ncu -i /tmp/ncu/vgemm.ncu-rep --page raw --csv > gpurun_out/vgemm_umma_r02_raw.csv 2>&1
ls -la /tmp/ncu
