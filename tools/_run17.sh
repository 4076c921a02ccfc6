set -x
mkdir -p gpurun_out
: > gpurun_out/gemm_min_tiles.jsonl
for mt in 1 150 300 600 1200 2400; do
  RT1B200_GEMM_PERSISTENT_MIN_TILES=$mt timeout -s KILL 600 python bench.py --reps 3 --no-e2e 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(json.dumps({'min_tiles': $mt, 'ms_per_step': d['ms_per_step'], 'value': d['value'], 'clocks': d['clocks']}))" >> gpurun_out/gemm_min_tiles.jsonl
done
cat gpurun_out/gemm_min_tiles.jsonl
