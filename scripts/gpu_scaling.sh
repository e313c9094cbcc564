#!/bin/bash
# strong-scaling sweep of bench.py on one box: bash scripts/gpu_scaling.sh "1 2 4 8" "pfsp:21:2 nq:17"
set -u
NS=${1:-"1 2 4 8"}
WL=${2:-"pfsp:21:2 nq:17"}
mkdir -p gpurun_out
port=29600
for w in $WL; do
  tag=$(echo $w | tr ':' '_')
  for n in $NS; do
    port=$((port+1))
    if [ "$n" = "1" ]; then
      python bench.py --gpus 1 --steps 3 --warmup 3 --workload $w --no-cpu-baseline > gpurun_out/scale_${tag}_n$n.json 2> gpurun_out/scale_${tag}_n$n.err
    else
      python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --gpus $n --steps 3 --warmup 3 --workload $w --no-cpu-baseline > gpurun_out/scale_${tag}_n$n.json 2> gpurun_out/scale_${tag}_n$n.err
    fi
    python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/scale_${tag}_n$n.json").read().strip().splitlines()[-1])
    print("$w n=$n value %.4g nodes/s ms/step %.1f e2e %.4g steals %s moved %s" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["config"].get("steals_per_step"), d["config"].get("nodes_moved_per_step")))
except Exception as e:
    print("$w n=$n FAILED", e); print(open("gpurun_out/scale_${tag}_n$n.err").read()[-1500:])
PY
  done
done
