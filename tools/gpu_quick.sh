#!/bin/bash
# quick check: kernel tests + short bench (+ optional launch list).  usage: gpu_quick.sh [launches]
mkdir -p gpurun_out
timeout -s KILL 900 python -m pytest tests/test_kernels_gpu.py tests/test_cnn_kernels_gpu.py -m gpu -q --timeout 300 -p no:cacheprovider -x > gpurun_out/pytest_quick.log 2>&1
echo "pytest exit $?"; grep -E "passed|failed" gpurun_out/pytest_quick.log | tail -2; grep -E "^FAILED|^ERROR|Error|error" gpurun_out/pytest_quick.log | head -20
timeout -s KILL 900 python bench.py --steps 20 --warmup 5 --no-e2e > gpurun_out/bench_quick.log 2>&1
echo "bench exit $?"; tail -1 gpurun_out/bench_quick.log | cut -c1-260
if [ "$1" = "launches" ]; then
timeout -s KILL 600 python tools/profile_step.py > gpurun_out/plain.log 2>&1 &&
timeout -s KILL 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
    --log-file gpurun_out/launches.csv python tools/profile_step.py > gpurun_out/ncu_launches.log 2>&1
echo "exit $?"; python tools/agg_launches.py gpurun_out/launches.csv 40 2>&1 | tee gpurun_out/launches_agg.txt
fi
