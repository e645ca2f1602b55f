set -x
timeout 600 python -m pytest tests/test_batch_gpu.py -m gpu -q -x 2>&1 | tail -2
timeout 300 python tests/dev_batch_time.py 120 8 2>&1 | tail -1
timeout 300 python tests/dev_phase_profile.py 600 > gpurun_out/phase_b1_n600.txt 2>&1; head -40 gpurun_out/phase_b1_n600.txt
