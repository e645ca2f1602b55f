set -x
timeout 300 python tests/dev_phase_profile.py 40 > gpurun_out/phase_b1.txt 2>&1; grep -E "att |frames_ms" gpurun_out/phase_b1.txt
