set -x
timeout 300 python tests/dev_frame_time.py 2>&1 | tail -1
timeout 300 python tests/dev_batch_time.py 120 8 2>&1 | tail -1
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_batch_gpu.py -m gpu -q -x 2>&1 | tail -3
