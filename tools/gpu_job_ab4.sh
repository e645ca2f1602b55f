# A/B 4: producer pacing sweep (piece size x gap) on the batch-1 kernel and on the 16-sequence kernel; HEAD library for reference
set -x
P=qwen3-tts.cpp_b200
mkdir -p gpurun_out
cp $P/libq3t_b200.so /tmp/new.so
cp $P/build/libq3t_base.so $P/libq3t_b200.so
( echo base; Q3T_AB_LENIENT=1 timeout 200 python tools/dev/dev_frame_time.py 2>&1 | tail -1 ) > gpurun_out/ab4.txt
cp /tmp/new.so $P/libq3t_b200.so
( echo new
  Q3T_PROD_PACE=0 timeout 200 python tools/dev/dev_frame_time.py 2>&1 | tail -1
  for pc in 1 2 4 8; do for pp in 1 100 200 300; do Q3T_PROD_PIECE=$pc Q3T_PROD_PACE=$pp timeout 200 python tools/dev/dev_frame_time.py 2>&1 | tail -1; done; done
  Q3T_PROD_PIECE=2 Q3T_PROD_PACE=50 timeout 200 python tools/dev/dev_frame_time.py 2>&1 | tail -1
  Q3T_PROD_PIECE=4 Q3T_PROD_PACE=150 timeout 200 python tools/dev/dev_frame_time.py 2>&1 | tail -1
  Q3T_PROD_PIECE=4 Q3T_PROD_PACE=200 Q3T_PROD_QUIET=1500 timeout 200 python tools/dev/dev_frame_time.py 2>&1 | tail -1
  Q3T_WTYPE=q8_0 Q3T_PROD_PACE=0 timeout 200 python tools/dev/dev_frame_time.py 2>&1 | tail -1
  Q3T_WTYPE=q8_0 timeout 200 python tools/dev/dev_frame_time.py 2>&1 | tail -1
) >> gpurun_out/ab4.txt
cat gpurun_out/ab4.txt
( for cfg in "0 4" "1 4" "100 4" "200 4" "300 4" "100 2" "200 8"; do set -- $cfg; Q3T_PROD_PACE=$1 Q3T_PROD_PIECE=$2 timeout 300 python tools/dev/dev_batch_time.py 120 16 2>&1 | grep batch | sed "s/^/pace $1 piece $2: /"; done ) > gpurun_out/batch_time_ab4.txt
cat gpurun_out/batch_time_ab4.txt
timeout 900 python -m pytest tests/test_batch_gpu.py -m gpu -q -x 2>&1 | tail -4
