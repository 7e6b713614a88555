# round 2, GPU call T: pruned plans derived from the unpruned plan — correctness, rebuild timing A/B, default bench line
cd /root/repo; mkdir -p gpurun_out
python -m pytest tests/test_gpu_tiles.py tests/test_gpu_parity.py tests/test_gpu_hostio.py -q -m gpu -x 2>&1 | grep -v "^$" | tail -25 > gpurun_out/t_tests.log
python -m pytest tests/test_gpu_fullsize.py -q -m gpu -x -k "c5 or c3" 2>&1 | grep -v "^$" | tail -6 >> gpurun_out/t_tests.log
python tools/bench_nlist.py eydpd > gpurun_out/t_nlist_derived.log 2>&1
AZB_NO_DERIVED_PLANS=1 python tools/bench_nlist.py eydpd > gpurun_out/t_nlist_hashed.log 2>&1
python bench.py > gpurun_out/t_bench.json 2> gpurun_out/t_bench.err; echo "rc=$?" >> gpurun_out/t_bench.err
tail -n 12 gpurun_out/t_tests.log; cat gpurun_out/t_nlist_derived.log gpurun_out/t_nlist_hashed.log; tail -2 gpurun_out/t_bench.err; head -c 600 gpurun_out/t_bench.json
