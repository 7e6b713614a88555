python -m pytest tests/test_gpu_tiles.py tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r1i_pytest.log
run() { timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-extras "$@" 2> gpurun_out/r1i_err.log | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$*', d['roofline']['kernel_ms'], d['value'], d['roofline']['frac'])" >> gpurun_out/r1i_sweep.log 2>&1; }
for tpp in 2 4 8; do run --workload dpd --tpp $tpp; done
run --workload dpd --tpp 4 --no-tiles
