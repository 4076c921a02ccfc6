#!/bin/bash
mkdir -p gpurun_out
echo "== pytest gpu (all single-GPU tests)"
timeout -s KILL 900 python -m pytest tests -m "gpu and not multigpu" -q --timeout 300 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "exit $?"; tail -25 gpurun_out/pytest_gpu.log
echo "== launch list of one training step"
timeout -s KILL 600 python tools/profile_step.py > gpurun_out/profile_plain.log 2>&1 && \
timeout -s KILL 1500 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches_step.csv python tools/profile_step.py > gpurun_out/profile_ncu.log 2>&1
echo "exit $?"; tail -3 gpurun_out/profile_ncu.log; wc -l gpurun_out/launches_step.csv
