#!/bin/bash
# final single-GPU verification of a round: tests, smoke, bench lines, reference arm, ncu evidence
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/final_tests.log 2>&1; tail -2 gpurun_out/final_tests.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python bench.py > gpurun_out/final_bench_ta021_lb2.json 2> gpurun_out/final_bench.err; tail -c 300 gpurun_out/final_bench.err
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/final_bench_reference.json 2>&1
python bench.py --workload nq:17 --ref-seconds 5 > gpurun_out/final_bench_nq17.json 2>/dev/null
python bench.py --workload pfsp:22:0 --no-cpu-baseline > gpurun_out/final_bench_ta022_lb1d.json 2>/dev/null
python bench.py --workload pfsp:14:1 --ref-seconds 3 > gpurun_out/final_bench_ta014_lb1.json 2>/dev/null
python bench.py --lb2-pairloop --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/final_bench_ta021_lb2_pairloop.json 2>/dev/null
python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/prof_bench_plain.json 2> gpurun_out/prof_bench_plain.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches_bench_final.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/prof_bench_ncu.log 2>&1
for spec in "pfsp 21 2 2097152:700" "pfsp 22 0 2097152:150" "nq 17 4194304:900"; do
  args=${spec%%:*}; skip=${spec##*:}
  tag=$(echo $args | cut -d' ' -f1-3 | tr ' ' '_')
  python scripts/profile_run.py $args > gpurun_out/plain_$tag.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:expand -s $skip -c 1 -f -o gpurun_out/prof4_$tag python scripts/profile_run.py $args > gpurun_out/ncu4_$tag.log 2>&1
done
ls gpurun_out | grep "final_\|prof4" 
