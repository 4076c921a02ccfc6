set -x
mkdir -p gpurun_out
RT1B200_TRAIN_GRAPH_STRICT=1 timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/test_gpu.log 2>&1; tail -12 gpurun_out/test_gpu.log | cut -c1-300
timeout 600 python tools/bench_dwconv.py --out gpurun_out/dwconv_bench2.json 2>&1 | tail -2 | cut -c1-400
RT1B200_TRAIN_GRAPH_STRICT=1 timeout 300 python bench.py --steps 10 --warmup 4 > gpurun_out/bench_lite1.json 2> gpurun_out/bench_lite1.err; tail -3 gpurun_out/bench_lite1.err | cut -c1-300; cut -c1-330 gpurun_out/bench_lite1.json
RT1B200_GEMM_LITE=0 timeout 300 python bench.py --steps 10 --warmup 4 --no-e2e > gpurun_out/bench_lite0.json 2>/dev/null; cut -c1-200 gpurun_out/bench_lite0.json
timeout 300 python tools/profile_step.py > gpurun_out/plain.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches.csv python tools/profile_step.py > gpurun_out/ncu_list.log 2>&1; tail -2 gpurun_out/ncu_list.log
python tools/agg_launches.py gpurun_out/launches.csv 45
