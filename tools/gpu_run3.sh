#!/bin/bash
mkdir -p gpurun_out
echo "== pytest cnn kernels"
timeout -s KILL 900 python -m pytest tests/test_cnn_kernels_gpu.py -m gpu -q --timeout 300 -p no:cacheprovider > gpurun_out/pytest_cnn.log 2>&1
echo "exit $?"; grep -E "passed|failed" gpurun_out/pytest_cnn.log | tail -3; grep -E "^FAILED|^ERROR" gpurun_out/pytest_cnn.log | head -40
echo "== pytest other gpu"
timeout -s KILL 900 python -m pytest tests -m "gpu and not multigpu" -q --timeout 300 -p no:cacheprovider --deselect tests/test_cnn_kernels_gpu.py > gpurun_out/pytest_gpu.log 2>&1
echo "exit $?"; tail -5 gpurun_out/pytest_gpu.log
echo "== smoke"
timeout -s KILL 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1
echo "exit $?"; tail -5 gpurun_out/smoke.log
echo "== bench ours"
timeout -s KILL 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_ours.log 2>&1
echo "exit $?"; tail -3 gpurun_out/bench_ours.log
