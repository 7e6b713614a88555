# round 2, GPU call V: lean plan builder (row-wise groups, warp-sorted runs + rank merge) — correctness, timings
cd /root/repo; mkdir -p gpurun_out
python -m pytest tests/test_gpu_tiles.py tests/test_gpu_parity.py tests/test_gpu_hostio.py tests/test_gpu_nlist.py tests/test_golden_fixtures.py -q -m gpu -x 2>&1 | grep -v "^$" | tail -25 > gpurun_out/v_tests.log
python -m pytest tests/test_gpu_fullsize.py -q -m gpu -x 2>&1 | grep -v "^$" | tail -6 >> gpurun_out/v_tests.log
for w in dpd plj4m eydpd; do python tools/bench_nlist.py $w; done > gpurun_out/v_nlist.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 160 --csv --log-file gpurun_out/v_launches_eydpd.csv python tools/bench_nlist.py eydpd > gpurun_out/v_ncu1.log 2>&1
tail -n 12 gpurun_out/v_tests.log; cat gpurun_out/v_nlist.log; grep "tiles_\|nlist_bcast" gpurun_out/v_launches_eydpd.csv | awk -F'","' '{print $5, $NF}' | sort | uniq -c | sort -rn | head -12
