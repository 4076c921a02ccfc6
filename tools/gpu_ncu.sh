#!/bin/bash
# ncu --set full captures of selected launches of ONE training step (tools/profile_step.py brackets it with
# cudaProfilerStart/Stop).  usage: gpu_ncu.sh "tag|regex|ordinal" ...
mkdir -p gpurun_out
timeout -s KILL 600 python tools/profile_step.py > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
for spec in "$@"; do
  IFS='|' read -r tag rx ord <<< "$spec"
  timeout -s KILL 900 ncu --set full --clock-control none --import-source on --profile-from-start off \
      --kernel-name-base demangled -k "regex:$rx" -s "$ord" -c 1 -f -o gpurun_out/ncu_$tag \
      python tools/profile_step.py > gpurun_out/ncu_$tag.log 2>&1
  echo "$tag exit $?"
done
ls -la gpurun_out/*.ncu-rep
