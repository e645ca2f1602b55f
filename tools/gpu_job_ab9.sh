# Q8_0 slices as contiguous ring ranges + warp-aggregated sampler histogram: parity, frame time, config 3
set -x
mkdir -p gpurun_out
( Q3T_WTYPE=q8_0 timeout 200 python tools/dev/dev_frame_time.py 2>&1 | tail -1
  Q3T_WTYPE=q8_0 Q3T_PROD_PACE=0 timeout 200 python tools/dev/dev_frame_time.py 2>&1 | tail -1
  Q3T_WTYPE=q8_0 Q3T_L2_PIN=20 timeout 200 python tools/dev/dev_frame_time.py 2>&1 | tail -1
  timeout 200 python tools/dev/dev_frame_time.py 2>&1 | tail -1 ) > gpurun_out/ab9.txt 2>&1
cat gpurun_out/ab9.txt
timeout 1200 python -m pytest tests/test_parity_gpu.py tests/test_sampling_gpu.py tests/test_batch_gpu.py tests/test_longform_gpu.py tests/test_variants_gpu.py -m gpu -q -x 2>&1 | tail -3
python bench.py --steps 2 --warmup 3 --wtype q8_0 --temperature 0.9 --top-k 50 --seed 7 --no-cpu-baseline --no-batch > gpurun_out/bench_ab9_config3.json 2> gpurun_out/bench_ab9_config3.err; echo "bench config3 rc=$?"
python bench.py --steps 2 --warmup 3 --wtype q8_0 --no-cpu-baseline --no-batch > gpurun_out/bench_ab9_q8.json 2> gpurun_out/bench_ab9_q8.err; echo "bench q8 rc=$?"
