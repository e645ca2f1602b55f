set -x
timeout 120 python tests/dev_voc_only.py 2>&1 | tail -2
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/voc_launches.csv python tests/dev_voc_only.py > gpurun_out/ncu_voc.log 2>&1
tail -2 gpurun_out/ncu_voc.log
