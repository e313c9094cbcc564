#!/bin/bash
# multi-GPU checkpoint: usage scripts/gpu_r2_multi.sh <N> [tests]
set -u
N=${1:-2}
mkdir -p gpurun_out/r2
if [ "${2:-}" = "tests" ]; then
  python -m pytest tests/test_gpu_engine.py -x -q -m gpu -k "multi_gpu or infinite_upper" 2>&1 | tail -5
fi
for w in pfsp:21:2 nq:17; do
  out=gpurun_out/r2/bench_n${N}_${w//:/_}
  BB_TRACE_PHASES=${BB_TRACE_PHASES:-} BB_TRACE_TEAM=${BB_TRACE_TEAM:-} python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 \
    bench.py --gpus $N --steps 5 --warmup 3 --workload $w > $out.json 2> $out.err
  echo "== $w N=$N rc=$?"
  python - <<PY
import json
try:
    d = json.loads([l for l in open("$out.json").read().splitlines() if l.startswith("{")][-1])
    c = d["config"]
    print({k: d[k] for k in ("value", "ms_per_step")}, "e2e", d["e2e"]["value"], "warm", c["warm_up_s_per_step"], "busy", [round(x, 2) for x in c["gpu_busy_fraction_per_rank"]], "steals", c["steals_per_step"], d.get("parity_checks"))
except Exception as ex:
    print("no line:", ex); print(open("$out.err").read()[-3000:])
PY
done
