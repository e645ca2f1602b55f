set -x
timeout 300 python tests/dev_phase_profile.py 40 8 > gpurun_out/phase_b8.txt 2>&1; head -34 gpurun_out/phase_b8.txt
