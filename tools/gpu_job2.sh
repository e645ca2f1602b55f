set -x
timeout 300 python tests/dev_phase_profile.py 40 8 > gpurun_out/phase_b8.txt 2>&1; tail -45 gpurun_out/phase_b8.txt
timeout 300 python tests/dev_batch_time.py 120 1,2,4,8 2>&1 | tail -5
python tests/dev_ncu_target.py 24 > gpurun_out/plain2.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:lm_batch_kernel -s 30 -c 1 -o gpurun_out/lm_batch_r1e python tests/dev_ncu_target.py 24 > gpurun_out/ncu2.log 2>&1
tail -3 gpurun_out/ncu2.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:vgemm_umma -c 80 -o gpurun_out/vgemm_umma_r1e python tests/dev_ncu_target.py 24 > gpurun_out/ncu3.log 2>&1
tail -3 gpurun_out/ncu3.log
