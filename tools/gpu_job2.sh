set -x
timeout 200 python tests/dev_wall_vs_events.py 320 smi 2>&1 | tail -10
python bench.py --steps 1 --warmup 1 --frames 24 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:lm_kernel -s 40 -c 1 -o gpurun_out/lm_kernel_r1c python bench.py --steps 1 --warmup 1 --frames 24 --no-cpu-baseline > gpurun_out/ncu2.log 2>&1
tail -3 gpurun_out/ncu2.log
