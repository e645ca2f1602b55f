# A/B 6: L2 cache-policy hints (evict_last on the first N weight slices of the code predictor, 4 per layer; evict_first on everything read once per frame)
set -x
mkdir -p gpurun_out
( for n in 0 6 7 8 9 10 11 12; do Q3T_L2_PIN=$n timeout 200 python tools/dev/dev_frame_time.py 2>&1 | tail -1; done
  for n in 0 12 16 20; do Q3T_WTYPE=q8_0 Q3T_L2_PIN=$n timeout 200 python tools/dev/dev_frame_time.py 2>&1 | tail -1; done
  for p in 100 150 250; do Q3T_L2_PIN=8 Q3T_PROD_PACE=$p timeout 200 python tools/dev/dev_frame_time.py 2>&1 | tail -1; done
) > gpurun_out/ab6b.txt 2>&1
cat gpurun_out/ab6b.txt
