#!/bin/bash
# GPU checkpoint: parity tests, short bench lines, ncu launch list + full captures (run via gpurun)
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/gpu_tests.log
tail -3 gpurun_out/gpu_tests.log
python bench.py --steps 1 --warmup 1 --ref-seconds 5 > gpurun_out/bench_ta021_lb2.json 2> gpurun_out/bench_ta021_lb2.err
python bench.py --steps 2 --warmup 3 --workload pfsp:14:1 --M 50000 --ref-seconds 5 > gpurun_out/bench_ta014_lb1.json 2> gpurun_out/bench_ta014_lb1.err
python bench.py --steps 1 --warmup 1 --workload nq:17 --ref-seconds 5 > gpurun_out/bench_nq17.json 2> gpurun_out/bench_nq17.err
python bench.py --steps 1 --warmup 1 --workload pfsp:22:0 --no-cpu-baseline > gpurun_out/bench_ta022_lb1d.json 2> gpurun_out/bench_ta022_lb1d.err
for spec in "pfsp 29 2" "pfsp 30 0" "nq 15"; do
  tag=$(echo $spec | tr ' ' '_')
  python scripts/profile_run.py $spec > gpurun_out/plain_$tag.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:expand -s 9 -c 2 -f -o gpurun_out/prof_$tag python scripts/profile_run.py $spec > gpurun_out/ncu_$tag.log 2>&1
done
ls -la gpurun_out
