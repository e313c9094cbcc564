#!/bin/bash
set -u
mkdir -p gpurun_out
python scripts/profile_run.py pfsp 29 2 > gpurun_out/plain_lb2s.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:expand_lb2 -s 10 -c 2 -f -o gpurun_out/prof_lb2_staged python scripts/profile_run.py pfsp 29 2 > gpurun_out/ncu_lb2s.log 2>&1
for ka in 2 4 6; do python - <<PY
import sys, time
sys.path.insert(0, '.')
from __graft_entry__ import load_package
bb = load_package()
ins = bb.Instance(inst=21)
res, per = bb.pfsp_search(ins, 2, M=1 << 20, lb2_stage_a=$ka)
p = per[0]
print("ta021 KA=$ka", res.tree, "kernel %.1f ms pairs/ev %.1f" % (res.kernel_ms_max, p.lb2_pairs / max(p.lb2_evals, 1)), flush=True)
PY
done
