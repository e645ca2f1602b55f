set -x
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_batch_gpu.py tests/test_longform_gpu.py -m gpu -q -x -k "generate or batched or step_vs or long_context" 2>&1 | tail -4
timeout 300 python tests/dev_phase_profile.py 40 > gpurun_out/phase_b1.txt 2>&1; grep -E "att |frames_ms" gpurun_out/phase_b1.txt
timeout 300 python tests/dev_frame_time.py 2>&1 | tail -1
timeout 300 python tests/dev_batch_time.py 120 8 2>&1 | tail -1
