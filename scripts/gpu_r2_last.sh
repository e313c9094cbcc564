#!/bin/bash
# last single-GPU pass of round 2: whole GPU suite, smoke, default bench line, N-Queens bench line + ncu of its big launch
set -u
out=gpurun_out/r2last; mkdir -p $out
timeout 900 python -m pytest tests -m gpu -x -q > $out/tests.log 2>&1; tail -2 $out/tests.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 300 python bench.py > $out/bench_ta021_lb2.json 2> $out/bench.err; tail -c 300 $out/bench.err
timeout 300 python bench.py --workload nq:17 --ref-seconds 5 > $out/bench_nq17.json 2>/dev/null
for f in $out/bench_*.json; do python - "$f" <<'PY'
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1].split("/")[-1], d.get("value"), d.get("ms_per_step"), (d.get("e2e") or {}).get("value"), (d.get("roofline") or {}).get("frac"))
PY
done
timeout 200 python scripts/profile_run.py nq 17 8388608 > $out/plain_nq.log 2>&1 &&
timeout 400 ncu --set full --clock-control none --import-source on -k regex:expand -s 7 -c 1 -f -o $out/prof_n_nq_17 python scripts/profile_run.py nq 17 8388608 > $out/ncu_n_nq.log 2>&1
tail -1 $out/plain_nq.log
