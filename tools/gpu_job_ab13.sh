# A/B 13: per-CTA candidate by warp shuffles (tree) vs serial scan (nocand); batched defaults; parity
set -x
P=qwen3-tts.cpp_b200
mkdir -p gpurun_out
cp $P/libq3t_b200.so /tmp/new.so
( cp $P/build/libq3t_nocand.so $P/libq3t_b200.so
  echo "== serial candidate scan"; timeout 200 python tools/dev/dev_frame_time.py 2>&1 | tail -1; timeout 200 python tools/dev/dev_frame_time.py 2>&1 | tail -1
  cp /tmp/new.so $P/libq3t_b200.so
  echo "== tree"; timeout 200 python tools/dev/dev_frame_time.py 2>&1 | tail -1; timeout 200 python tools/dev/dev_frame_time.py 2>&1 | tail -1
  timeout 300 python tools/dev/dev_batch_time.py 120 16,64 2>&1 | grep "^batch" ) > gpurun_out/ab13.txt 2>&1
grep -v "^+" gpurun_out/ab13.txt
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_batch_gpu.py tests/test_sampling_gpu.py -m gpu -q -x 2>&1 | tail -2
