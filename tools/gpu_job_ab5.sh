# A/B 5 (one box): HEAD vs tree vs tree with per-layer K/V publication (v1) vs tree with the old FP64 chains (v2) vs both (v3);
# the same for the 16-sequence kernel; Q8_0 frame time
set -x
P=qwen3-tts.cpp_b200
mkdir -p gpurun_out
cp $P/libq3t_b200.so /tmp/new.so
run() { cp $1 $P/libq3t_b200.so; echo "== $2"; Q3T_AB_LENIENT=1 timeout 200 python tools/dev/dev_frame_time.py 2>&1 | tail -1; Q3T_AB_LENIENT=1 timeout 200 python tools/dev/dev_frame_time.py 2>&1 | tail -1; }
( run $P/build/libq3t_base.so base
  run /tmp/new.so new
  run $P/build/libq3t_v1.so "v1 per-layer publication"
  run $P/build/libq3t_v2.so "v2 old FP64 chains"
  run $P/build/libq3t_v3.so "v3 both old"
  cp /tmp/new.so $P/libq3t_b200.so
  echo "== new q8_0"; Q3T_WTYPE=q8_0 timeout 200 python tools/dev/dev_frame_time.py 2>&1 | tail -1
) > gpurun_out/ab5.txt 2>&1
cat gpurun_out/ab5.txt
brun() { cp $1 $P/libq3t_b200.so; timeout 300 python tools/dev/dev_batch_time.py 120 16 2>&1 | grep batch | sed "s/^/$2: /"; }
( brun /tmp/new.so new; brun $P/build/libq3t_v1.so v1; brun $P/build/libq3t_v2.so v2; brun $P/build/libq3t_v3.so v3 ) > gpurun_out/batch_time_ab5.txt 2>&1
cat gpurun_out/batch_time_ab5.txt
cp /tmp/new.so $P/libq3t_b200.so
