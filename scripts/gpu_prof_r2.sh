#!/bin/bash
# ncu --set full of one mid-search launch of each dominant kernel (usage: scripts/gpu_prof_r2.sh <tag>)
set -u
tag=${1:-a}
mkdir -p gpurun_out/r2
for spec in "pfsp 21 2 8388608:200" "nq 17 8388608:8" "pfsp 22 0 4194304:80"; do
  args=${spec%%:*}; skip=${spec##*:}
  name=$(echo $args | cut -d' ' -f1-3 | tr ' ' '_')
  python scripts/profile_run.py $args > gpurun_out/r2/plain_${name}.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:expand -s $skip -c 1 -f -o gpurun_out/r2/prof_${tag}_${name} python scripts/profile_run.py $args > gpurun_out/r2/ncu_${tag}_${name}.log 2>&1
  tail -1 gpurun_out/r2/plain_${name}.log
done
ls -la gpurun_out/r2 | grep ncu-rep
