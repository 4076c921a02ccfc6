set -x
mkdir -p gpurun_out
timeout -s KILL 900 python -m pytest tests/test_cnn_kernels_gpu.py -m gpu -q 2>&1 | tail -3
timeout -s KILL 600 python bench.py --reps 3 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(json.dumps({'ms_per_step': d['ms_per_step'], 'value': d['value'], 'e2e': d['e2e']['value']}))"
timeout -s KILL 600 python tools/bench_dwconv.py --out gpurun_out/dwconv_bench.json 2>&1 | tail -1 | cut -c1-400
