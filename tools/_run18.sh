set -x
mkdir -p gpurun_out
: > gpurun_out/gemm_dbg.jsonl
for dbg in 0 2 1 3; do
  RT1B200_GEMM_DBG=$dbg timeout -s KILL 600 python bench.py --reps 3 --no-e2e 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(json.dumps({'gemm_dbg': $dbg, 'ms_per_step': d['ms_per_step'], 'clocks': d['clocks']}))" >> gpurun_out/gemm_dbg.jsonl
done
cat gpurun_out/gemm_dbg.jsonl
