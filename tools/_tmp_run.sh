set -x
timeout -s KILL 600 python -m pytest tests/test_cnn_kernels_gpu.py -m gpu -q 2>&1 | tail -2
cp rt1b200/lib/librt1b200.so /tmp/keep.so
for rep in 1 2; do for v in old new; do cp rt1b200/lib/variants/$v.so rt1b200/lib/librt1b200.so; timeout -s KILL 600 python bench.py --reps 3 --no-e2e 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$v', d['ms_per_step'])"; done; done
cp /tmp/keep.so rt1b200/lib/librt1b200.so
