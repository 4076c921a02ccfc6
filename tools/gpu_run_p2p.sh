#!/bin/bash
# multi-GPU: real-process P2P tests + sweep. usage: gpu_run_p2p.sh NGPU
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_$N.txt 2>&1
echo "== p2p worker on $N GPUs (with sweep)"
timeout -s KILL 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1 --master-port 29631 tests/_p2p_worker.py --sweep > gpurun_out/p2p_worker_$N.log 2>&1
echo "exit $?"; tail -30 gpurun_out/p2p_worker_$N.log
echo "== bench ours $N GPUs"
timeout -s KILL 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1 --master-port 29632 bench.py --gpus $N --steps 8 --warmup 3 > gpurun_out/bench_ours_$N.log 2>&1
echo "exit $?"; tail -3 gpurun_out/bench_ours_$N.log
echo "== bench reference $N GPUs"
timeout -s KILL 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1 --master-port 29633 bench.py --impl reference --gpus $N --steps 8 --warmup 3 > gpurun_out/bench_ref_$N.log 2>&1
echo "exit $?"; tail -3 gpurun_out/bench_ref_$N.log
