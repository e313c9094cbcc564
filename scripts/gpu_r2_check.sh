#!/bin/bash
# round-2 GPU checkpoint: the -m gpu suite, then bench lines of the four workloads (no CPU baseline leg)
set -u
mkdir -p gpurun_out/r2
python -m pytest tests -x -q -m gpu 2>&1 | tail -15 > gpurun_out/r2/pytest_gpu.txt
cat gpurun_out/r2/pytest_gpu.txt
for w in pfsp:21:2 nq:17 pfsp:22:0 pfsp:14:1; do
  python bench.py --workload $w --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2/bench_${w//:/_}.json 2> gpurun_out/r2/bench_${w//:/_}.err
  echo "== $w rc=$?"; python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r2/bench_${w//:/_}.json").read().strip().splitlines()[-1])
    print({k: d[k] for k in ("value", "ms_per_step")}, d["e2e"]["value"], d["roofline"]["frac"], d["config"].get("device_iterations_per_step"))
except Exception as ex:
    print("no line:", ex); print(open("gpurun_out/r2/bench_${w//:/_}.err").read()[-2000:])
PY
done
