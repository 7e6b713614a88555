# round 2, GPU call Q: rows in place (no compaction) — correctness + rebuild timings
cd /root/repo; mkdir -p gpurun_out
python -m pytest tests/test_gpu_nlist.py tests/test_gpu_md.py tests/test_gpu_tiles.py tests/test_gpu_hostio.py tests/test_zz_no_alloc.py -q -m gpu -x 2>&1 | grep -v "^$" | tail -15 > gpurun_out/q_tests.log
python -m pytest tests/test_gpu_fullsize.py -q -m gpu -x 2>&1 | grep -v "^$" | tail -6 >> gpurun_out/q_tests.log
for w in dpd plj4m eydpd; do python tools/bench_nlist.py $w; done > gpurun_out/q_nlist.log 2>&1
AZB_NLIST_PACKED=1 python tools/bench_nlist.py eydpd > gpurun_out/q_nlist_packed.log 2>&1
tail -n 12 gpurun_out/q_tests.log; cat gpurun_out/q_nlist.log gpurun_out/q_nlist_packed.log
