#!/bin/bash
# N-Queens: tests of the kernels, then a sweep over the number of levels enumerated in registers (BB_NQ_TAIL) and M
timeout 400 python -m pytest tests/test_gpu_engine.py tests/test_gpu_pins.py tests/test_gpu_share.py -x -q -m gpu -k "nq or queens or share or generations or several or process or pool_agrees" 2>&1 | tail -6
for K in ${TAILS:-7 8 9}; do
  for M in ${MS:-2097152 8388608}; do
    BB_NQ_TAIL=$K timeout 120 python bench.py --workload nq:17 --M $M --steps 3 --warmup 2 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('tail', $K, 'M', $M, round(d['value']/1e9,1), round(d['ms_per_step'],2), d['config'].get('device_iterations_per_step'), round(d['e2e']['value']/1e9,1))"
  done
done
