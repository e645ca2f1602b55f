# A/B 7: L2 prefetch one layer ahead of the ring copies, on top of the pinned code-predictor slices and the paced producer
set -x
mkdir -p gpurun_out
( timeout 200 python tools/dev/dev_frame_time.py 2>&1 | tail -1
  Q3T_L2_AHEAD=1 timeout 200 python tools/dev/dev_frame_time.py 2>&1 | tail -1
  Q3T_L2_AHEAD=1 Q3T_L2_PIN=6 timeout 200 python tools/dev/dev_frame_time.py 2>&1 | tail -1
  Q3T_L2_AHEAD=1 Q3T_PROD_PACE=0 timeout 200 python tools/dev/dev_frame_time.py 2>&1 | tail -1
  Q3T_L2_AHEAD=1 Q3T_PROD_PACE=50 timeout 200 python tools/dev/dev_frame_time.py 2>&1 | tail -1
  Q3T_WTYPE=q8_0 timeout 200 python tools/dev/dev_frame_time.py 2>&1 | tail -1
  Q3T_WTYPE=q8_0 Q3T_L2_AHEAD=1 timeout 200 python tools/dev/dev_frame_time.py 2>&1 | tail -1
  timeout 200 python tools/dev/dev_frame_time.py 2>&1 | tail -1 ) > gpurun_out/ab7.txt 2>&1
cat gpurun_out/ab7.txt
timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -q -x -k "generate or step or prefill" 2>&1 | tail -2
