set -x
mkdir -p gpurun_out
RT1B200_TRAIN_GRAPH_STRICT=1 timeout 900 python -m pytest tests -m gpu -q > gpurun_out/test_gpu.log 2>&1; tail -6 gpurun_out/test_gpu.log | cut -c1-300
RT1B200_TRAIN_GRAPH_STRICT=1 timeout 300 python bench.py --steps 10 --warmup 4 > gpurun_out/bench_1gpu.json 2> gpurun_out/bench_1gpu.err; tail -3 gpurun_out/bench_1gpu.err | cut -c1-300; cut -c1-330 gpurun_out/bench_1gpu.json
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 | cut -c1-300
