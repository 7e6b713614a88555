cd /root/repo; mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_golden_fixtures.py tests/test_gpu_fullsize.py tests/test_gpu_hostio.py tests/test_gpu_md.py -x -q -m gpu -k "dpd or c5 or golden or host" 2>&1 | tail -4 > gpurun_out/s14t_tests.log
run() { python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-extras "$@" 2>> gpurun_out/s14t_err.log | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$*', d['ms_per_step'], d['roofline']['kernel_ms'], d['value'], d['roofline']['frac'], 'e2e', d['e2e']['ms_per_step'])" >> gpurun_out/s14t_sweep.log 2>&1; }
run --workload dpd
run --workload dpd
run --workload eydpd
cat gpurun_out/s14t_tests.log gpurun_out/s14t_sweep.log; tail -5 gpurun_out/s14t_err.log
