#!/bin/bash
# end-of-round 1-GPU record: full GPU test suite, smoke, both bench arms (with e2e), launch list
mkdir -p gpurun_out
timeout -s KILL 1200 python -m pytest tests -m gpu -q --timeout 600 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?"; grep -E "passed|failed" gpurun_out/pytest_gpu.log | tail -2; grep -E "^FAILED|^ERROR" gpurun_out/pytest_gpu.log | head
timeout -s KILL 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout -s KILL 900 python bench.py --steps 30 --warmup 5 2>&1 | tail -1 | tee gpurun_out/bench_ours_final.json | cut -c1-300
timeout -s KILL 900 python bench.py --impl reference --steps 10 --warmup 3 2>&1 | tail -1 | tee gpurun_out/bench_ref_final.json | cut -c1-200
timeout -s KILL 600 python tools/bench_policy.py --impl ours 2>/dev/null | tail -1 | tee gpurun_out/policy_ours.json
timeout -s KILL 900 python bench.py --model scaled --steps 10 --warmup 3 --no-e2e 2>&1 | tail -1 | tee gpurun_out/bench_scaled_1.json | cut -c1-200
