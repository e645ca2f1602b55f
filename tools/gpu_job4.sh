set -x
timeout 300 python tests/dev_phase_profile.py 40 8 > gpurun_out/phase_b8_e.txt 2>&1; grep -E "E |norm|frames_ms" gpurun_out/phase_b8_e.txt | head -30
for d1 in 0 150 400 800; do echo "POLL_D1=$d1"; Q3T_POLL_D1=$d1 timeout 200 python tests/dev_batch_time.py 60 1,8 2>&1 | grep "^batch"; done
