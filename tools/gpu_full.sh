#!/bin/bash
# One-GPU round trip: all GPU tests, smoke, both bench arms, then the per-launch list of one training step.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/gpu.txt 2>&1
echo "== pytest -m gpu"
timeout -s KILL 1200 python -m pytest tests -m gpu -q --timeout 600 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "exit $?"; grep -E "passed|failed" gpurun_out/pytest_gpu.log | tail -3; grep -E "^FAILED|^ERROR" gpurun_out/pytest_gpu.log | head -40
echo "== smoke"
timeout -s KILL 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1
echo "exit $?"; tail -2 gpurun_out/smoke.log
echo "== bench ours"
timeout -s KILL 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_ours.log 2>&1
echo "exit $?"; tail -1 gpurun_out/bench_ours.log
if [ "$1" != "noref" ]; then
echo "== bench reference"
timeout -s KILL 900 python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/bench_ref.log 2>&1
echo "exit $?"; tail -1 gpurun_out/bench_ref.log
fi
echo "== launch list (one step)"
timeout -s KILL 600 python tools/profile_step.py > gpurun_out/plain.log 2>&1 &&
timeout -s KILL 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
    --log-file gpurun_out/launches.csv python tools/profile_step.py > gpurun_out/ncu_launches.log 2>&1
echo "exit $?"; python tools/agg_launches.py gpurun_out/launches.csv 45 2>&1 | tee gpurun_out/launches_agg.txt
