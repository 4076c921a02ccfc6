set -x
mkdir -p gpurun_out
timeout -s KILL 300 python tools/profile_p2p_emulated.py 2>&1 | tail -2
timeout -s KILL 600 ncu --set full --clock-control none --import-source on --profile-from-start off --kernel-name-base demangled -k "regex:p2p_allreduce_adamw_kernel" -c 1 -f -o gpurun_out/ncu_p2p_emulated python tools/profile_p2p_emulated.py > gpurun_out/ncu_p2p_emulated.log 2>&1; tail -3 gpurun_out/ncu_p2p_emulated.log
for impl in ours; do timeout -s KILL 600 python tools/bench_policy.py --impl $impl > gpurun_out/policy_$impl.json 2>gpurun_out/policy_$impl.err; tail -1 gpurun_out/policy_$impl.json | cut -c1-900; done
