cd /root/repo; mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4) > gpurun_out/s14r_tests.log
run() { python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-extras "$@" 2>> gpurun_out/s14r_err.log | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$*', d['ms_per_step'], d['roofline']['kernel_ms'], d['value'], d['roofline']['frac'], 'e2e', d['e2e']['ms_per_step'])" >> gpurun_out/s14r_sweep.log 2>&1; }
run --workload dpd
run --workload plj4m
run --workload patchy
run --workload melt
run --workload hertz
run --workload colloid
run --workload eydpd
cat gpurun_out/s14r_tests.log gpurun_out/s14r_sweep.log; tail -5 gpurun_out/s14r_err.log
