set -x
mkdir -p gpurun_out
nvidia-smi -L
RT1B200_TRAIN_GRAPH_STRICT=1 timeout 900 python -m pytest tests/test_trainer_gpu.py tests/test_p2p_gpu.py tests/test_data_parallel.py -x -q -m gpu > gpurun_out/test_mgpu.log 2>&1; tail -15 gpurun_out/test_mgpu.log | cut -c1-600
RT1B200_TRAIN_GRAPH_STRICT=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29701 bench.py --gpus 2 --steps 10 --warmup 4 > gpurun_out/bench_2gpu.json 2> gpurun_out/bench_2gpu.err; tail -5 gpurun_out/bench_2gpu.err | cut -c1-400; cat gpurun_out/bench_2gpu.json
