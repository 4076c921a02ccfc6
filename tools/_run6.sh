set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_cnn_kernels_gpu.py -q -x -k "full_encoder" 2>&1 | tail -4 | cut -c1-200
RT1B200_GEMM_BALANCED=0 timeout 600 python -m pytest tests/test_cnn_kernels_gpu.py -q -x -k "full_encoder" 2>&1 | tail -4 | cut -c1-200
RT1B200_TRAIN_GRAPH_STRICT=1 timeout 900 python -m pytest tests -m gpu -q > gpurun_out/test_gpu.log 2>&1; tail -8 gpurun_out/test_gpu.log | cut -c1-300
timeout 600 python tools/bench_dwconv.py --out gpurun_out/dwconv_bench3.json --layers b2,b3-4,b6-7,b14-17 2>&1 | grep -o '"layer": "[^"]*"\|"fwd_act_us": [0-9.]*\|"wgrad_act_us": [0-9.]*\|"fwd_us": [0-9.]*' | paste - - - - 
RT1B200_TRAIN_GRAPH_STRICT=1 timeout 300 python bench.py --steps 10 --warmup 4 --no-e2e > gpurun_out/bench_a.json 2>/dev/null; cut -c1-200 gpurun_out/bench_a.json
RT1B200_GEMM_BALANCED=0 RT1B200_GEMM_LITE=0 timeout 300 python bench.py --steps 10 --warmup 4 --no-e2e > gpurun_out/bench_b.json 2>/dev/null; cut -c1-200 gpurun_out/bench_b.json
