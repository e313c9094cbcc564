#!/bin/bash
# round-2 single-GPU evidence: tests, smoke, bench lines (default with CPU baseline, reference arm, other workloads),
# ncu launch list of the bench command, ncu --set full of one mid-search launch of each dominant kernel
#   usage: scripts/gpu_r2_final.sh <tag>
set -u
tag=${1:-z}
out=gpurun_out/r2final; mkdir -p $out
timeout 900 python -m pytest tests -m gpu -x -q > $out/tests.log 2>&1; tail -2 $out/tests.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 300 python bench.py > $out/bench_ta021_lb2.json 2> $out/bench.err; tail -c 300 $out/bench.err
timeout 300 python bench.py --impl reference --steps 1 --warmup 0 > $out/bench_reference.json 2>&1
timeout 300 python bench.py --workload nq:17 --ref-seconds 5 > $out/bench_nq17.json 2>/dev/null
timeout 300 python bench.py --workload pfsp:22:0 --no-cpu-baseline > $out/bench_ta022_lb1d.json 2>/dev/null
timeout 300 python bench.py --workload pfsp:14:1 --ref-seconds 3 > $out/bench_ta014_lb1.json 2>/dev/null
for f in $out/bench_*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1].split("/")[-1], d.get("value"), d.get("ms_per_step"), (d.get("e2e") or {}).get("value"), (d.get("roofline") or {}).get("frac"))
except Exception as ex:
    print(sys.argv[1], "no line:", ex)
PY
done
timeout 300 python bench.py --steps 1 --warmup 1 --no-cpu-baseline > $out/prof_bench_plain.json 2> $out/prof_bench_plain.err &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $out/launches_bench.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > $out/prof_bench_ncu.log 2>&1
for spec in "pfsp 21 2 8388608:200" "nq 17 8388608:8" "pfsp 22 0 4194304:80"; do
  args=${spec%%:*}; skip=${spec##*:}
  name=$(echo $args | tr ' ' '_')
  timeout 200 python scripts/profile_run.py $args > $out/plain_${name}.log 2>&1 &&
  timeout 400 ncu --set full --clock-control none --import-source on -k regex:expand -s $skip -c 1 -f -o $out/prof_${tag}_${name} python scripts/profile_run.py $args > $out/ncu_${tag}_${name}.log 2>&1
  tail -1 $out/plain_${name}.log
done
ls -la $out | grep ncu-rep
