# N-GPU evidence run (usage: gpurun --gpus N -- bash tools/gpu_run_multi.sh N [sweep]): multi-GPU tests, both bench arms,
# the scaled variant, training on disk-backed episodes, optionally the fused-exchange sweep
set -x
mkdir -p gpurun_out
N=${1:-8}
nproc; free -g | head -2
RT1B200_TRAIN_GRAPH_STRICT=1 timeout -s KILL 900 python -m pytest tests/test_p2p_gpu.py tests/test_data_parallel.py -q -m gpu > gpurun_out/test_mgpu_$N.log 2>&1; tail -4 gpurun_out/test_mgpu_$N.log | cut -c1-400
run() { tag=$1; shift; timeout -s KILL 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1 --master-port 29632 bench.py --gpus $N "$@" > gpurun_out/bench_${tag}_$N.json 2> gpurun_out/bench_${tag}_$N.err; echo "$tag exit $?"; grep -a "^{" gpurun_out/bench_${tag}_$N.json | tail -1 | cut -c1-420; }
RT1B200_TRAIN_GRAPH_STRICT=1 run ours --steps 20 --warmup 4
run ref --impl reference --steps 10 --warmup 3
RT1B200_TRAIN_GRAPH_STRICT=1 run ours_scaled --model scaled --steps 20 --warmup 4
timeout -s KILL 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1 --master-port 29633 tools/bench_input_pipeline.py --workers 4 --episodes 96 --steps 40 --out gpurun_out/input_pipeline_$N.json > gpurun_out/input_pipeline_$N.log 2>&1; echo "input exit $?"; grep -a "^{" gpurun_out/input_pipeline_$N.log | tail -1 | cut -c1-600
if [ "$2" = "sweep" ]; then timeout -s KILL 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1 --master-port 29631 tests/_p2p_worker.py --sweep > gpurun_out/p2p_worker_$N.log 2>&1; echo "sweep exit $?"; ls gpurun_out/p2p_sweep_w$N.json; fi
