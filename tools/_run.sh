cd /root/repo; mkdir -p gpurun_out
run() { python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-extras "$@" 2>> gpurun_out/r2d_err.log | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$*', d['ms_per_step'], d['roofline']['kernel_ms'], d['value'], d['roofline']['frac'], d['e2e']['value'])" >> gpurun_out/r2d_sweep.log 2>&1; }
run --workload plj4m
AZB_BANK_ROTATE=1 run --workload plj4m
run --workload plj4m
AZB_BANK_ROTATE=1 run --workload plj4m
AZB_BANK_ROTATE=1 run --workload melt
run --workload melt
AZB_BANK_ROTATE=1 run --workload patchy
run --workload patchy
