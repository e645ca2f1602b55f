set -x
timeout 900 python -m pytest tests/test_longform_gpu.py tests/test_batch_gpu.py -m gpu -q -x -k "long_context or batched" 2>&1 | tail -2
timeout 300 python tests/dev_frame_time.py 2>&1 | tail -1
timeout 300 python tests/dev_phase_profile.py 600 > gpurun_out/phase_b1_n600.txt 2>&1; grep -E "frames_ms|att K|att P ready|attention done" gpurun_out/phase_b1_n600.txt
