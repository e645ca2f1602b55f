# A/B 11: L2 prefetch of every CTA's local candidate's embedding row (greedy), side build vs tree
set -x
P=qwen3-tts.cpp_b200
mkdir -p gpurun_out
cp $P/libq3t_b200.so /tmp/new.so
( echo "== tree"; timeout 200 python tools/dev/dev_frame_time.py 2>&1 | tail -1; timeout 200 python tools/dev/dev_frame_time.py 2>&1 | tail -1
  cp $P/build/libq3t_pf.so $P/libq3t_b200.so
  echo "== candidate-row prefetch"; timeout 200 python tools/dev/dev_frame_time.py 2>&1 | tail -1; timeout 200 python tools/dev/dev_frame_time.py 2>&1 | tail -1
  timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -q -x -k "generate" 2>&1 | tail -2
  cp /tmp/new.so $P/libq3t_b200.so ) > gpurun_out/ab11.txt 2>&1
grep -v "^+" gpurun_out/ab11.txt
