set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py -x -q -k "attention" > gpurun_out/test_attn.log 2>&1; tail -25 gpurun_out/test_attn.log | cut -c1-300
RT1B200_TRAIN_GRAPH_STRICT=1 timeout 900 python -m pytest tests -m gpu -q > gpurun_out/test_gpu.log 2>&1; tail -25 gpurun_out/test_gpu.log | cut -c1-300
RT1B200_TRAIN_GRAPH_STRICT=1 timeout 300 python bench.py --steps 10 --warmup 4 > gpurun_out/bench_attn_tc.json 2> gpurun_out/bench_attn_tc.err; tail -3 gpurun_out/bench_attn_tc.err | cut -c1-300; cut -c1-400 gpurun_out/bench_attn_tc.json
RT1B200_ATTN_TC=0 timeout 300 python bench.py --steps 10 --warmup 4 --no-e2e > gpurun_out/bench_attn_cc.json 2>/dev/null; cut -c1-200 gpurun_out/bench_attn_cc.json
