set -x
timeout 600 python -m pytest tests/test_vocoder_gpu.py -m gpu -q -x 2>&1 | tail -3
timeout 120 python tests/dev_voc_only.py 2>&1 | tail -2
timeout 120 python tests/dev_voc_only.py 4096 2>&1 | tail -2
