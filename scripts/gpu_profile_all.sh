#!/bin/bash
# launch list of the bench command + ncu --set full of the three expand kernels (run via gpurun, one GPU)
set -u
mkdir -p gpurun_out
python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/prof_bench_plain.json 2> gpurun_out/prof_bench_plain.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 9000 --csv --log-file gpurun_out/launches_bench.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/prof_bench_ncu.log 2>&1
for spec in "pfsp 21 2:1500" "pfsp 22 0:300" "nq 17:3000"; do
  args=${spec%%:*}; skip=${spec##*:}
  tag=$(echo $args | tr ' ' '_')
  python scripts/profile_run.py $args > gpurun_out/plain_$tag.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:expand -s $skip -c 2 -f -o gpurun_out/prof2_$tag python scripts/profile_run.py $args > gpurun_out/ncu2_$tag.log 2>&1
done
ls -la gpurun_out | tail -12
