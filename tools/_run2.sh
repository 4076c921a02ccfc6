set -x
mkdir -p gpurun_out
RT1B200_TRAIN_GRAPH_STRICT=1 timeout 600 python -m pytest tests/test_trainer_gpu.py -x -q 2>&1 | tail -12 | cut -c1-400
timeout 300 python tools/diag_graph.py 2>&1 | grep -v Warning | tail -8 | cut -c1-400
RT1B200_TRAIN_GRAPH_STRICT=1 timeout 300 python bench.py --steps 10 --warmup 4 > gpurun_out/bench_graph.json 2> gpurun_out/bench_graph.err; tail -5 gpurun_out/bench_graph.err; cat gpurun_out/bench_graph.json
RT1B200_TRAIN_GRAPH=0 timeout 300 python bench.py --steps 10 --warmup 4 --no-e2e > gpurun_out/bench_eager.json 2> gpurun_out/bench_eager.err; cat gpurun_out/bench_eager.json
timeout 300 python tools/profile_step.py > gpurun_out/plain.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches.csv python tools/profile_step.py > gpurun_out/ncu_list.log 2>&1; tail -3 gpurun_out/ncu_list.log
python tools/agg_launches.py gpurun_out/launches.csv 60
