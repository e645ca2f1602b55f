set -x
timeout 300 python tests/dev_batch64.py 320 2>&1 | grep -v "backend" | tail -6
timeout 300 python tests/dev_phase_profile.py 40 > gpurun_out/phase_b1.txt 2>&1; head -48 gpurun_out/phase_b1.txt
