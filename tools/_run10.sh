set -x
mkdir -p gpurun_out
RT1B200_TRAIN_GRAPH_STRICT=1 timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/test_gpu.log 2>&1; tail -8 gpurun_out/test_gpu.log | cut -c1-300
for impl in ours raster; do timeout 600 python tools/bench_policy.py --impl $impl > gpurun_out/policy_$impl.json 2>gpurun_out/policy_$impl.err; tail -1 gpurun_out/policy_$impl.json | cut -c1-700; done
export RT1B200_TRAIN_GRAPH=0
bash tools/gpu_ncu.sh "dw_fwd_k5s1|dw_tile_kernel<5, 1, 0,|2" "dw_bwd_k5s1|dw_tile_kernel<5, 1, 1,|2" "dw_wgrad_k5s1|dw_tile_kernel<5, 1, 2,|2" "dw_fwd_k3s1|dw_tile_kernel<3, 1, 0,|1" "attn_tc_fwd|attn_tc_fwd_kernel|0" "attn_tc_bwd|attn_tc_bwd_kernel|0" "se_bn_bwd_reduce|se_bn_bwd_reduce_kernel|8" "gemm_persistent|gemm_tcgen05_persistent_kernel|20"
