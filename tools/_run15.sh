set -x
mkdir -p gpurun_out
RT1B200_TRAIN_GRAPH_STRICT=1 timeout -s KILL 1200 python -m pytest tests -m gpu -q > gpurun_out/test_gpu.log 2>&1; tail -6 gpurun_out/test_gpu.log | cut -c1-300
timeout -s KILL 600 python bench.py --reps 3 > gpurun_out/bench_ours_1gpu.json 2>gpurun_out/bench_ours_1gpu.err; tail -1 gpurun_out/bench_ours_1gpu.json | cut -c1-1500
