# batched prefill + final defaults: batch tests, batch-64 timing, bench line
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_batch_gpu.py tests/test_pipeline_gpu.py -m gpu -q -x 2>&1 | tail -4
Q3T_BATCH_TRACE=1 timeout 300 python tools/dev/dev_batch_time.py 120 16,64 > gpurun_out/batch_time_ab8.txt 2>&1; grep -E "batch|generate_batch" gpurun_out/batch_time_ab8.txt | tail -8
python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_ab8.json 2> gpurun_out/bench_ab8.err; echo "bench rc=$?"
