# 8-GPU run: multi-GPU correctness, fused exchange sweep, both bench arms (RT-1 and the scaled variant)
set -x
mkdir -p gpurun_out
N=${1:-8}
nvidia-smi topo -m > gpurun_out/topo_$N.txt 2>&1
RT1B200_TRAIN_GRAPH_STRICT=1 timeout 900 python -m pytest tests/test_p2p_gpu.py tests/test_data_parallel.py -q -m gpu > gpurun_out/test_mgpu_$N.log 2>&1; tail -6 gpurun_out/test_mgpu_$N.log | cut -c1-400
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1 --master-port 29631 tests/_p2p_worker.py --sweep > gpurun_out/p2p_worker_$N.log 2>&1
echo "worker exit $?"; grep -v "^\[W\|^W1\|^\*\*\*\|OMP_NUM" gpurun_out/p2p_worker_$N.log | tail -30 | cut -c1-500
run() { tag=$1; shift; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1 --master-port 29632 bench.py --gpus $N "$@" > gpurun_out/bench_${tag}_$N.json 2> gpurun_out/bench_${tag}_$N.err; echo "$tag exit $?"; tail -1 gpurun_out/bench_${tag}_$N.json | cut -c1-420; }
RT1B200_TRAIN_GRAPH_STRICT=1 run ours --steps 20 --warmup 4
run ref --impl reference --steps 10 --warmup 3
RT1B200_TRAIN_GRAPH_STRICT=1 run ours_scaled --model scaled --steps 20 --warmup 4
run ref_scaled --impl reference --model scaled --steps 10 --warmup 3
run ref_bf16 --impl reference --ref-precision bf16 --steps 10 --warmup 3 --no-e2e
