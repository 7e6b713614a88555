cd /root/repo; mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/r2a_pytest.log
run() { python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras "$@" 2>> gpurun_out/r2a_err.log | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$*', d['ms_per_step'], d['roofline']['kernel_ms'], d['value'], d['roofline']['frac'], d['e2e']['value'])" >> gpurun_out/r2a_sweep.log 2>&1; }
run --workload plj4m
AZB_NO_PACKED=1 run --workload plj4m
run --workload dpd
run --workload eydpd
run --workload melt
