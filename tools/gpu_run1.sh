#!/bin/bash
# First GPU bring-up: kernel numerics, smoke, both bench arms. Each stage in its own process with its own timeout.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
echo "== pytest gpu kernels" 
timeout -s KILL 900 python -m pytest tests/test_kernels_gpu.py -m gpu -q --timeout 180 -p no:cacheprovider > gpurun_out/pytest_kernels.log 2>&1
echo "exit $?"; tail -40 gpurun_out/pytest_kernels.log
echo "== smoke"
timeout -s KILL 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1
echo "exit $?"; tail -15 gpurun_out/smoke.log
echo "== bench ours"
timeout -s KILL 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_ours.log 2>&1
echo "exit $?"; tail -5 gpurun_out/bench_ours.log
echo "== bench reference"
timeout -s KILL 900 python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/bench_ref.log 2>&1
echo "exit $?"; tail -5 gpurun_out/bench_ref.log
