cd /root/repo; mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -q -m gpu 2>&1 | grep -v "^$" | tail -15 > gpurun_out/i_noalloc_alone.log
python -m pytest tests -q -m gpu -x 2>&1 | grep -v "^$" | tail -15 > gpurun_out/i_tests.log
for w in dpd plj4m eydpd; do python tools/bench_nlist.py $w; done > gpurun_out/i_nlist.log 2>&1
AZB_NO_FUSED_PLANS=1 python tools/bench_nlist.py eydpd >> gpurun_out/i_nlist.log 2>&1
AZB_NO_FUSED_PLANS=1 python tools/bench_nlist.py dpd >> gpurun_out/i_nlist.log 2>&1
tail -5 gpurun_out/i_noalloc_alone.log; tail -6 gpurun_out/i_tests.log; cat gpurun_out/i_nlist.log
