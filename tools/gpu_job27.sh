set -x
timeout 300 python tests/dev_phase_profile.py 600 > gpurun_out/phase_b1_n600.txt 2>&1; grep -E "frames_ms|att K|K slice" gpurun_out/phase_b1_n600.txt
