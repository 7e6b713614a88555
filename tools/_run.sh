cd /root/repo; mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r1y_pytest.log
for w in dpd plj4m; do timeout 300 python tools/bench_nlist.py $w >> gpurun_out/r1y_nlist.log 2>&1; done
run() { python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras "$@" 2>> gpurun_out/r1y_err.log | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$*', d['roofline']['kernel_ms'], d['value'], d['roofline']['frac'], d['e2e']['value'])" >> gpurun_out/r1y_sweep.log 2>&1; }
run --workload plj4m
run --workload dpd
run --workload patchy
