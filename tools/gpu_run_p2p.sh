#!/bin/bash
# multi-GPU: emulated-rank tests, real-process P2P tests + sweep, then both bench arms. usage: gpu_run_p2p.sh NGPU [nobench]
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_$N.txt 2>&1
echo "== emulated p2p tests"
timeout -s KILL 600 python -m pytest tests/test_p2p_gpu.py -m "gpu and not multigpu" -q -p no:cacheprovider 2>&1 | tail -3
echo "== p2p worker on $N GPUs (with sweep)"
timeout -s KILL 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1 --master-port 29631 tests/_p2p_worker.py --sweep > gpurun_out/p2p_worker_$N.log 2>&1
echo "exit $?"; grep -v "^\[W\|^W1\|^\*\*\*\|OMP_NUM" gpurun_out/p2p_worker_$N.log | tail -40 | cut -c1-400
[ "$2" = "nobench" ] && exit 0
echo "== bench ours $N GPUs"
timeout -s KILL 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1 --master-port 29632 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_ours_$N.log 2>&1
echo "exit $?"; tail -1 gpurun_out/bench_ours_$N.log
echo "== bench reference $N GPUs"
timeout -s KILL 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1 --master-port 29633 bench.py --impl reference --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_ref_$N.log 2>&1
echo "exit $?"; tail -1 gpurun_out/bench_ref_$N.log
