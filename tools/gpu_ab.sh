# A/B of a side build (qwen3-tts.cpp_b200/build/libq3t_hoist.so) against the tree's library on one box
set -x
P=qwen3-tts.cpp_b200
timeout 300 python tools/dev/dev_frame_time.py 2>&1 | tail -1
timeout 300 python tools/dev/dev_frame_time.py 2>&1 | tail -1
timeout 300 python tools/dev/dev_batch_time.py 120 8 2>&1 | tail -1
cp $P/build/libq3t_hoist.so $P/libq3t_b200.so
echo "=== side build"
timeout 300 python tools/dev/dev_frame_time.py 2>&1 | tail -1
timeout 300 python tools/dev/dev_batch_time.py 120 8 2>&1 | tail -1
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_batch_gpu.py tests/test_longform_gpu.py -m gpu -q -x -k "generate or step_vs or prefill or batched or long_context or legacy" 2>&1 | tail -3
