python -m pytest tests/test_gpu_tiles.py tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r1e_pytest.log
run() { python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras "$@" 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$*', d['roofline']['kernel_ms'], d['value'], d['roofline']['frac'])" >> gpurun_out/r1e_sweep.log 2>&1; }
for tpp in 2 4 8; do run --workload plj4m --tpp $tpp; done
run --workload plj4m --tpp 4 --tile-size 256
run --workload plj4m --tpp 4 --tile-size 64
run --workload plj4m --no-tiles
run --workload plj4m --tpp 4 --no-virial
run --workload dpd
