set -x
mkdir -p gpurun_out
RT1B200_TRAIN_GRAPH_STRICT=1 timeout -s KILL 1200 python -m pytest tests -m gpu -q > gpurun_out/test_gpu.log 2>&1; tail -3 gpurun_out/test_gpu.log | cut -c1-300
timeout -s KILL 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1 | cut -c1-300
timeout -s KILL 600 python bench.py > gpurun_out/bench_ours_1gpu.json 2>/dev/null; tail -1 gpurun_out/bench_ours_1gpu.json | cut -c1-1400
export RT1B200_TRAIN_GRAPH=0
timeout -s KILL 300 python tools/profile_step.py > gpurun_out/plain.log 2>&1 && timeout -s KILL 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches.csv python tools/profile_step.py > gpurun_out/ncu_list.log 2>&1
python tools/agg_launches.py gpurun_out/launches.csv 70 > gpurun_out/step_launches.txt; head -6 gpurun_out/step_launches.txt
unset RT1B200_TRAIN_GRAPH
for impl in ours graphed; do timeout -s KILL 600 python tools/bench_policy.py --impl $impl > gpurun_out/policy_$impl.json 2>gpurun_out/policy_$impl.err; tail -1 gpurun_out/policy_$impl.json | cut -c1-500; done
