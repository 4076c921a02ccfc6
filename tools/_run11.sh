set -x
mkdir -p gpurun_out
RT1B200_TRAIN_GRAPH_STRICT=1 timeout 1200 python -m pytest tests -m gpu -q -s > gpurun_out/test_gpu.log 2>&1; grep -a "convergence\]" gpurun_out/test_gpu.log | cut -c1-400; tail -8 gpurun_out/test_gpu.log | cut -c1-300
timeout 900 python bench.py > gpurun_out/bench_ours_tail1.json 2>gpurun_out/bench_ours_tail1.err; tail -1 gpurun_out/bench_ours_tail1.json | cut -c1-900
RT1B200_BN_TAIL=0 timeout 900 python bench.py > gpurun_out/bench_ours_tail0.json 2>gpurun_out/bench_ours_tail0.err; tail -1 gpurun_out/bench_ours_tail0.json | cut -c1-400
timeout 900 python bench.py --impl reference > gpurun_out/bench_ref_1gpu.json 2>gpurun_out/bench_ref_1gpu.err; tail -1 gpurun_out/bench_ref_1gpu.json | cut -c1-500
export RT1B200_TRAIN_GRAPH=0
timeout 300 python tools/profile_step.py > gpurun_out/plain.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches.csv python tools/profile_step.py > gpurun_out/ncu_list.log 2>&1; tail -2 gpurun_out/ncu_list.log
python tools/agg_launches.py gpurun_out/launches.csv 70 > gpurun_out/step_launches.txt; head -40 gpurun_out/step_launches.txt
bash tools/gpu_ncu.sh "dw_fwd_k5s1|dw_tile_kernel<.int.5, .int.1, .int.0,|2" "dw_bwd_k5s1|dw_tile_kernel<.int.5, .int.1, .int.1,|2" "dw_wgrad_k5s1|dw_tile_kernel<.int.5, .int.1, .int.2,|2" "dw_fwd_k3s1|dw_tile_kernel<.int.3, .int.1, .int.0,|1" "bn_act_bwd_apply|bn_act_bwd_kernel<.bool.1|10" "se_bwd_a|se_bwd_a_kernel|20" "se_bwd_b|se_bwd_b_kernel|20"
