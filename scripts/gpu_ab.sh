#!/bin/bash
# A/B of an environment switch on the three headline workloads (usage: scripts/gpu_ab.sh VAR=1 [M_pfsp M_nq])
sw=$1; mp=${2:-2097152}; mq=${3:-4194304}
for w in nq:17 pfsp:22:0 pfsp:21:2; do
  M=$mp; [ $w = nq:17 ] && M=$mq
  for v in "" "$sw"; do
    for i in 1 2; do
      env $v timeout 200 python bench.py --workload $w --M $M --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$w', '[$v]', $M, round(d['value']/1e9,2), round(d['ms_per_step'],2), round(d['roofline']['avg_launch_us'],1), round(d['e2e']['value']/1e9,2))"
    done
  done
done
