# Round-2 job: smoke + GPU tests, then the measurement job (tools/gpu_profile_job_r02.sh)
set -x
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_r02.log 2>&1; echo "smoke rc=$?"
timeout 1700 python -m pytest tests -m gpu -q -x --durations=15 > gpurun_out/pytest_gpu_r02.log 2>&1; echo "pytest rc=$?"
tail -25 gpurun_out/pytest_gpu_r02.log
bash tools/gpu_profile_job_r02.sh
