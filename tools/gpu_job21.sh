set -x
for k in 1024 512 256; do echo "=== MT_K $k"; Q3T_VOC_MT_K=$k timeout 120 python tests/dev_voc_only.py 2>&1 | tail -2; done
python bench.py --steps 2 --warmup 3 --wtype q8_0 --temperature 0.9 --top-k 50 --seed 7 --no-cpu-baseline > gpurun_out/bench_r1g_config3.json 2> gpurun_out/bench_r1g_config3.err; echo "bench config3 rc=$?"; cat gpurun_out/bench_r1g_config3.json | cut -c1-900
