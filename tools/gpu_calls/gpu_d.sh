# round 2, GPU call D: pruned plans, tile cache, hostio; bench default (short)
cd /root/repo; mkdir -p gpurun_out
python -m pytest tests/test_gpu_tiles.py tests/test_gpu_hostio.py tests/test_gpu_parity.py tests/test_gpu_md.py tests/test_gpu_nlist.py -q -m gpu -x 2>&1 | grep -v "^$" | tail -30 > gpurun_out/d_tests.log
python -m pytest tests/test_gpu_fullsize.py -q -m gpu -x -k "c5 or c3" -s 2>&1 | grep -v "^$" | tail -12 >> gpurun_out/d_tests.log
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/d_bench.json 2> gpurun_out/d_bench.err
tail -8 gpurun_out/d_tests.log; tail -3 gpurun_out/d_bench.err; head -c 300 gpurun_out/d_bench.json
