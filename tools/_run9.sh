set -x
mkdir -p gpurun_out
RT1B200_TRAIN_GRAPH_STRICT=1 timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/test_gpu.log 2>&1; tail -8 gpurun_out/test_gpu.log | cut -c1-300
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 | cut -c1-300
timeout 600 python tools/bench_input_pipeline.py --out gpurun_out/input_pipeline.json 2>&1 | tail -2 | cut -c1-900
for impl in ours graphed reference raster; do timeout 600 python tools/bench_policy.py --impl $impl > gpurun_out/policy_$impl.json 2>gpurun_out/policy_$impl.err; tail -1 gpurun_out/policy_$impl.json | cut -c1-700; done
