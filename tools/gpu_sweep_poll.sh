for d0 in 0 150 300 450; do for d1 in 0 60; do Q3T_POLL_D0=$d0 Q3T_POLL_D1=$d1 timeout 120 python tests/dev_frame_time.py 60 2>&1 | tail -1; done; done
