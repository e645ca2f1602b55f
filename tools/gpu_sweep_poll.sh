for d1 in 0 -1; do Q3T_POLL_D0=200 Q3T_POLL_D1=$d1 timeout 120 python tools/dev/dev_frame_time.py 60 2>&1 | tail -1; done
for d1 in 0 -1; do Q3T_POLL_D0=200 Q3T_POLL_D1=$d1 timeout 120 python tools/dev/dev_frame_time.py 320 2>&1 | tail -1; done
