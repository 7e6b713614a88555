# round 2, GPU call R: direct-store emission A/B + in-place rows tests
cd /root/repo; mkdir -p gpurun_out
python -m pytest tests/test_gpu_nlist.py tests/test_gpu_md.py tests/test_gpu_tiles.py tests/test_gpu_hostio.py tests/test_zz_no_alloc.py -q -m gpu -x 2>&1 | grep -v "^$" | tail -25 > gpurun_out/r_tests.log
AZB_NLIST_BUFFERED=1 python -m pytest tests/test_gpu_nlist.py -q -m gpu -x 2>&1 | grep -v "^$" | tail -5 >> gpurun_out/r_tests.log
for w in dpd plj4m eydpd; do python tools/bench_nlist.py $w; done > gpurun_out/r_nlist_direct.log 2>&1
for w in dpd plj4m eydpd; do AZB_NLIST_BUFFERED=1 python tools/bench_nlist.py $w; done > gpurun_out/r_nlist_buffered.log 2>&1
tail -n 12 gpurun_out/r_tests.log; cat gpurun_out/r_nlist_direct.log gpurun_out/r_nlist_buffered.log
