set -x
timeout 300 python tests/dev_frame_time.py 2>&1 | tail -1
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_batch_gpu.py -m gpu -q -x 2>&1 | tail -3
timeout 300 python tests/dev_phase_profile.py 40 > gpurun_out/phase_b1.txt 2>&1; head -32 gpurun_out/phase_b1.txt
