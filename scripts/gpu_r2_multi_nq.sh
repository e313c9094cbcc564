#!/bin/bash
# N-Queens bench line on N GPUs (usage: scripts/gpu_r2_multi_nq.sh <N>)
N=${1:-2}; mkdir -p gpurun_out/r2
out=gpurun_out/r2/bench_n${N}_nq_17
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 5 --warmup 3 --workload nq:17 > $out.json 2> $out.err
echo "== nq:17 N=$N rc=$?"
python - <<PY
import json
d = json.loads([l for l in open("$out.json").read().splitlines() if l.startswith("{")][-1])
c = d["config"]
print({k: d[k] for k in ("value", "ms_per_step")}, "e2e", d["e2e"]["value"], "warm", c["warm_up_s_per_step"], "busy", [round(x, 2) for x in c["gpu_busy_fraction_per_rank"]], "steals", c["steals_per_step"], d.get("parity_checks"))
PY
