#!/bin/bash
# config 4 (policy latency, both arms), rasteriser, scaled-model bench on 1 GPU, full pytest
mkdir -p gpurun_out
for impl in ours reference raster; do
  timeout -s KILL 600 python tools/bench_policy.py --impl $impl 2> gpurun_out/policy_$impl.err | tail -1 | tee gpurun_out/policy_$impl.json
done
timeout -s KILL 900 python bench.py --model scaled --steps 10 --warmup 3 --no-e2e 2>&1 | tail -1 | tee gpurun_out/bench_scaled_1.json | cut -c1-400
timeout -s KILL 1200 python -m pytest tests -m gpu -q --timeout 600 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?"; grep -E "passed|failed" gpurun_out/pytest_gpu.log | tail -2; grep -E "^FAILED|^ERROR" gpurun_out/pytest_gpu.log | head
