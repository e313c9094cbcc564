#!/bin/bash
set -u
mkdir -p gpurun_out
for spec in "pfsp 21 2:1500" "pfsp 22 0:300"; do
  args=${spec%%:*}; skip=${spec##*:}
  tag=$(echo $args | tr ' ' '_')
  python scripts/profile_run.py $args > gpurun_out/plain_$tag.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:expand -s $skip -c 1 -f -o gpurun_out/prof3_$tag python scripts/profile_run.py $args > gpurun_out/ncu3_$tag.log 2>&1
done
python scripts/profile_run.py pfsp 21 2 4194304
python scripts/profile_run.py pfsp 21 2 2097152
python scripts/profile_run.py pfsp 22 0 4194304
