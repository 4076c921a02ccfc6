# multi-GPU check of the current tree: correctness tests, then our bench arm (RT-1 and scaled) and the reference arm
set -x
mkdir -p gpurun_out
N=${1:-2}
RT1B200_TRAIN_GRAPH_STRICT=1 timeout -s KILL 900 python -m pytest tests/test_p2p_gpu.py tests/test_data_parallel.py -q -m gpu > gpurun_out/test_mgpu_$N.log 2>&1; tail -4 gpurun_out/test_mgpu_$N.log | cut -c1-400
run() { tag=$1; shift; timeout -s KILL 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1 --master-port 29632 bench.py --gpus $N "$@" > gpurun_out/bench_${tag}_$N.json 2> gpurun_out/bench_${tag}_$N.err; echo "$tag exit $?"; grep -a "^{" gpurun_out/bench_${tag}_$N.json | tail -1 | cut -c1-1300; }
RT1B200_TRAIN_GRAPH_STRICT=1 run ours --steps 20 --warmup 4
RT1B200_TRAIN_GRAPH_STRICT=1 run ours_scaled --model scaled --steps 20 --warmup 4
if [ "$2" = "ref" ]; then run ref --impl reference --steps 10 --warmup 3; fi
