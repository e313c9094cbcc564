#!/bin/bash
# batch-size sweep of the three headline workloads (usage: scripts/gpu_sweep_m.sh)
for spec in "nq:17 1048576 2097152 4194304 8388608" "pfsp:22:0 524288 1048576 2097152 4194304" "pfsp:21:2 524288 1048576 2097152 4194304"; do
  set -- $spec; w=$1; shift
  for M in "$@"; do
    timeout 200 python bench.py --workload $w --M $M --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$w', $M, round(d['value']/1e9,2), round(d['ms_per_step'],2), round(d['roofline']['avg_launch_us'],1), round(d['e2e']['value']/1e9,2))"
  done
done
