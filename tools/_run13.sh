set -x
mkdir -p gpurun_out
RT1B200_TRAIN_GRAPH_STRICT=1 timeout -s KILL 1200 python -m pytest tests -m gpu -q > gpurun_out/test_gpu.log 2>&1; tail -6 gpurun_out/test_gpu.log | cut -c1-300
for v in 1 0; do RT1B200_PDL=$v timeout -s KILL 200 python tools/bench_pdl.py 2>&1 | tail -1; done | tee gpurun_out/pdl_microbench.jsonl
for impl in ours graphed; do timeout -s KILL 600 python tools/bench_policy.py --impl $impl > gpurun_out/policy_$impl.json 2>gpurun_out/policy_$impl.err; tail -1 gpurun_out/policy_$impl.json | cut -c1-600; done
