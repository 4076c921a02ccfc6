set -x
mkdir -p gpurun_out
# programmatic dependent launch is new: prove it terminates on one small step before anything long runs
if timeout -s KILL 240 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_pdl.log 2>&1; then tail -1 gpurun_out/smoke_pdl.log | cut -c1-200; else echo "SMOKE WITH PDL FAILED"; tail -5 gpurun_out/smoke_pdl.log | cut -c1-300; export RT1B200_PDL=0; fi
RT1B200_TRAIN_GRAPH_STRICT=1 timeout -s KILL 1200 python -m pytest tests -m gpu -q -s > gpurun_out/test_gpu.log 2>&1; grep -a "convergence\]" gpurun_out/test_gpu.log | cut -c1-500; tail -6 gpurun_out/test_gpu.log | cut -c1-300
for cfg in "1 1" "1 0" "0 0" "0 1"; do set -- $cfg; RT1B200_PDL=$1 RT1B200_BN_TAIL=$2 timeout -s KILL 600 python bench.py --reps 3 > gpurun_out/bench_ours_pdl$1_tail$2.json 2>gpurun_out/bench_ours_pdl$1_tail$2.err; tail -1 gpurun_out/bench_ours_pdl$1_tail$2.json | cut -c1-330; done
export RT1B200_TRAIN_GRAPH=0
timeout -s KILL 300 python tools/profile_step.py > gpurun_out/plain.log 2>&1 && timeout -s KILL 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches.csv python tools/profile_step.py > gpurun_out/ncu_list.log 2>&1; tail -2 gpurun_out/ncu_list.log
python tools/agg_launches.py gpurun_out/launches.csv 70 > gpurun_out/step_launches.txt; head -24 gpurun_out/step_launches.txt
bash tools/gpu_ncu.sh "dw_fwd_block6|dw_tile_kernel<.int.5, .int.1, .int.0,|0" "dw_bwd_block6|dw_tile_kernel<.int.5, .int.1, .int.1,|11" "dw_wgrad_block6|dw_tile_kernel<.int.5, .int.1, .int.2,|11" "bn_act_bwd_apply_big|bn_act_bwd_kernel<.bool.1|46"
