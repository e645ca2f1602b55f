# two GPUs: the default bench line (weak-scaling step + config 4 dealt over the ranks) under torchrun
set -x
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_r02_n2.json 2> gpurun_out/bench_r02_n2.err; echo "n2 rc=$?"; cut -c1-300 gpurun_out/bench_r02_n2.json; tail -3 gpurun_out/bench_r02_n2.err
timeout 600 python -m pytest tests/test_pipeline_gpu.py -m gpu -q -x 2>&1 | tail -3
