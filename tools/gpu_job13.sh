set -x
P=qwen3-tts.cpp_b200
for r in 4 8 1; do
  cp $P/build/libq3t_rep$r.so $P/libq3t_b200.so
  echo "=== REP $r"
  timeout 300 python tests/dev_frame_time.py 2>&1 | tail -1
  timeout 300 python tests/dev_batch_time.py 120 8 2>&1 | tail -1
done
cp $P/build/libq3t_rep4.so $P/libq3t_b200.so
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_batch_gpu.py -m gpu -q -x -k "generate or batched or step_vs" 2>&1 | tail -3
timeout 300 python tests/dev_phase_profile.py 40 8 > gpurun_out/phase_b8_rep4.txt 2>&1; head -30 gpurun_out/phase_b8_rep4.txt
