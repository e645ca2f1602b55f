set -x
timeout 300 tools/microbench/bin/umma_test > gpurun_out/umma_test.txt 2>&1; echo "umma rc=$?"; cat gpurun_out/umma_test.txt
timeout 600 python -m pytest tests/test_vocoder_gpu.py -m gpu -q -x 2>&1 | tail -5
timeout 300 python tests/dev_voc_only.py 2>&1 | tail -5
