for N in 14 15 16; do for K in 5 6 7 8 9; do
BB_NQ_TAIL=$K timeout 120 python bench.py --workload nq:$N --M 8388608 --steps 3 --warmup 2 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('N', $N, 'tail', $K, round(d['value']/1e9,1), round(d['ms_per_step'],3), d['config'].get('device_iterations_per_step'), round(d['e2e']['value']/1e9,1))"
done; done
BB_NQ_TAIL=9 timeout 120 python bench.py --workload nq:18 --M 8388608 --steps 2 --warmup 1 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('N 18 tail 9', round(d['value']/1e9,1), round(d['ms_per_step'],3), d['config'].get('device_iterations_per_step'))"
