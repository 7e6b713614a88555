cd /root/repo; mkdir -p gpurun_out
python -m pytest tests/test_gpu_hostio.py -x -q -m gpu 2>&1 | tail -8 > gpurun_out/s14e_tests.log
python tools/e2e_probe.py > gpurun_out/s14e_probe.log 2>&1
run() { python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-extras "$@" 2>> gpurun_out/s14e_err.log | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$*', d['ms_per_step'], d['roofline']['kernel_ms'], d['value'], d['roofline']['frac'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'])" >> gpurun_out/s14e_sweep.log 2>&1; }
run --workload dpd --e2e-chunks 8
run --workload dpd --e2e-chunks 12
run --workload dpd --e2e-chunks 16
run --workload dpd --e2e-chunks 24
run --workload plj4m --e2e-chunks 8
run --workload plj4m --e2e-chunks 16
run --workload plj4m --e2e-chunks 24
run --workload patchy --e2e-chunks 16
cat gpurun_out/s14e_tests.log gpurun_out/s14e_probe.log gpurun_out/s14e_sweep.log; tail -5 gpurun_out/s14e_err.log
