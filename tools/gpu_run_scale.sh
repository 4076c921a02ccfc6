#!/bin/bash
# N-GPU: P2P worker (tests + sweep) and our bench arm.  usage: gpu_run_scale.sh NGPU
N=${1:-8}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_$N.txt 2>&1
timeout -s KILL 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1 --master-port 29631 tests/_p2p_worker.py --sweep > gpurun_out/p2p_worker_$N.log 2>&1
echo "worker exit $?"; grep -E "^\{|^\[p2p\]|WORKER_OK|Error|error" gpurun_out/p2p_worker_$N.log | cut -c1-330 | tail -30
timeout -s KILL 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1 --master-port 29632 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_ours_$N.log 2>&1
echo "bench exit $?"; tail -1 gpurun_out/bench_ours_$N.log
