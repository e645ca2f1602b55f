set -x
Q3T_VOC_MT=2 timeout 120 tools/microbench/bin/umma_test > gpurun_out/umma_test_mt2.txt 2>&1; echo "umma mt2 rc=$?"; grep -E "FAIL|OK|CUDA" gpurun_out/umma_test_mt2.txt
timeout 120 tools/microbench/bin/umma_test > gpurun_out/umma_test.txt 2>&1; echo "umma rc=$?"; cat gpurun_out/umma_test.txt
timeout 600 python -m pytest tests/test_vocoder_gpu.py -m gpu -q -x 2>&1 | tail -3
timeout 120 python tests/dev_voc_only.py 2>&1 | tail -3
