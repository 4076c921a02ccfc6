set -x
mkdir -p gpurun_out
N=${1:-8}
run() { tag=$1; shift; timeout -s KILL 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1 --master-port 29632 bench.py --gpus $N "$@" > gpurun_out/bench_${tag}_$N.json 2> gpurun_out/bench_${tag}_$N.err; echo "$tag exit $?"; grep -a "^{" gpurun_out/bench_${tag}_$N.json | tail -1 | cut -c1-300; }
RT1B200_TRAIN_GRAPH_STRICT=1 run ours --steps 20 --warmup 4
RT1B200_TRAIN_GRAPH_STRICT=1 run ours_scaled --model scaled --steps 20 --warmup 4
