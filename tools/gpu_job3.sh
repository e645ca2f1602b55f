set -x
python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -4 gpurun_out/pytest_gpu.log
python tests/dev_ncu_target.py 24 > gpurun_out/plain2.log 2>&1 || exit 1
mkdir -p /tmp/ncu
timeout 600 ncu --set full --clock-control none -k regex:lm_batch_kernel -s 30 -c 1 -o /tmp/ncu/lm_batch python tests/dev_ncu_target.py 24 > gpurun_out/ncu2.log 2>&1
ncu -i /tmp/ncu/lm_batch.ncu-rep --page details > gpurun_out/lm_batch_r1e_details.txt 2>&1
ncu -i /tmp/ncu/lm_batch.ncu-rep --page raw --csv > gpurun_out/lm_batch_r1e_raw.csv 2>&1
timeout 900 ncu --set full --clock-control none -k regex:vgemm_umma -c 71 -o /tmp/ncu/vgemm python tests/dev_ncu_target.py 24 > gpurun_out/ncu3.log 2>&1
ncu -i /tmp/ncu/vgemm.ncu-rep --page raw --csv > gpurun_out/vgemm_umma_r1e_raw.csv 2>&1
ls -la /tmp/ncu gpurun_out
