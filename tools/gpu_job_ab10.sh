# bench line with the config-4 utterances decoded by 4 vocoder instances; 1 vs 4 instances
set -x
mkdir -p gpurun_out
python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_ab10.json 2> gpurun_out/bench_ab10.err; echo "bench rc=$?"
python bench.py --steps 1 --warmup 3 --no-cpu-baseline --decoders 1 > gpurun_out/bench_ab10_d1.json 2> gpurun_out/bench_ab10_d1.err; echo "bench rc=$?"
python bench.py --steps 1 --warmup 3 --no-cpu-baseline --decoders 2 > gpurun_out/bench_ab10_d2.json 2> gpurun_out/bench_ab10_d2.err; echo "bench rc=$?"
tail -3 gpurun_out/bench_ab10.err
