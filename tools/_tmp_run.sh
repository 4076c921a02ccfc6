set -x
mkdir -p gpurun_out
RT1B200_TRAIN_GRAPH_STRICT=1 timeout -s KILL 1200 python -m pytest tests -m gpu -q > gpurun_out/test_gpu.log 2>&1; tail -4 gpurun_out/test_gpu.log | cut -c1-300
timeout -s KILL 600 python bench.py --reps 3 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(json.dumps({'ms_per_step': d['ms_per_step'], 'value': d['value'], 'e2e': d['e2e']['value']}))"
