# final check of the committed tree: smoke, the whole GPU suite, the default bench line
set -x
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_final.log 2>&1; echo "smoke rc=$?"
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu_final.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/pytest_gpu_final.log
python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err; echo "bench rc=$?"
