# eight GPUs: the default bench line under torchrun (weak-scaling step + config 4 dealt over the ranks)
set -x
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 8 --steps 3 --warmup 3 > gpurun_out/bench_r02_n8.json 2> gpurun_out/bench_r02_n8.err; echo "n8 rc=$?"; cut -c1-200 gpurun_out/bench_r02_n8.json; tail -3 gpurun_out/bench_r02_n8.err
