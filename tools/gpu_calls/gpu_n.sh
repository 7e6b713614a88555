# round 2, GPU call N: broadcast neighbour-list walk — correctness, rebuild timings A/B, ncu of the rebuild kernels
cd /root/repo; mkdir -p gpurun_out
python -m pytest tests/test_gpu_nlist.py tests/test_gpu_md.py tests/test_gpu_tiles.py tests/test_zz_no_alloc.py -q -m gpu -x 2>&1 | grep -v "^$" | tail -15 > gpurun_out/n_tests.log
python -m pytest tests/test_gpu_fullsize.py -q -m gpu -x 2>&1 | grep -v "^$" | tail -6 >> gpurun_out/n_tests.log
for w in dpd plj4m eydpd; do python tools/bench_nlist.py $w; done > gpurun_out/n_nlist_bcast.log 2>&1
for w in dpd plj4m eydpd; do AZB_NLIST_LANEWISE=1 python tools/bench_nlist.py $w; done > gpurun_out/n_nlist_lanewise.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"nlist_bcast_kernel|tiles_build_flat|nlist_compact" --launch-skip 4 --launch-count 4 -o gpurun_out/n_rebuild -f python tools/bench_nlist.py eydpd > gpurun_out/n_ncu_rebuild.log 2>&1
python tools/ncu_summary.py gpurun_out/n_rebuild.ncu-rep gpurun_out/n_rebuild_summary.json "C5 8M rebuild kernels (broadcast walk, flat plan builder)" > gpurun_out/n_rebuild_summary.txt 2>&1
tail -n 8 gpurun_out/n_tests.log; cat gpurun_out/n_nlist_bcast.log gpurun_out/n_nlist_lanewise.log; ls -la gpurun_out/*.ncu-rep
