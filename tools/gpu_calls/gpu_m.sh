# round 2, GPU call M: flat tile-plan builder — correctness (tile / nlist / parity / C5 full size) and rebuild timings A/B
cd /root/repo; mkdir -p gpurun_out
python -m pytest tests/test_gpu_tiles.py tests/test_gpu_nlist.py tests/test_gpu_hostio.py -q -m gpu -x 2>&1 | grep -v "^$" | tail -15 > gpurun_out/m_tests.log
python -m pytest tests/test_gpu_fullsize.py -q -m gpu -x -k "c5 or c3" 2>&1 | grep -v "^$" | tail -6 >> gpurun_out/m_tests.log
for w in dpd plj4m eydpd; do python tools/bench_nlist.py $w; done > gpurun_out/m_nlist_flat.log 2>&1
for w in dpd plj4m eydpd; do AZB_TILES_ROWWISE=1 python tools/bench_nlist.py $w; done > gpurun_out/m_nlist_rowwise.log 2>&1
tail -n 8 gpurun_out/m_tests.log; cat gpurun_out/m_nlist_flat.log gpurun_out/m_nlist_rowwise.log
