set -x
mkdir -p gpurun_out
cp rt1b200/lib/librt1b200.so /tmp/lib_default.so
: > gpurun_out/pdl_modes.jsonl
for m in 0 1 2 3; do
  cp rt1b200/lib/variants/librt1b200_pdl$m.so rt1b200/lib/librt1b200.so
  echo "{\"mode\": $m}" >> gpurun_out/pdl_modes.jsonl
  timeout -s KILL 200 python tools/bench_pdl.py 2>&1 | tail -1 >> gpurun_out/pdl_modes.jsonl
  timeout -s KILL 600 python bench.py --reps 3 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(json.dumps({'ms_per_step': d['ms_per_step'], 'value': d['value'], 'e2e': d['e2e']['value'], 'clocks': d['clocks']}))" >> gpurun_out/pdl_modes.jsonl
done
cp /tmp/lib_default.so rt1b200/lib/librt1b200.so
cat gpurun_out/pdl_modes.jsonl
