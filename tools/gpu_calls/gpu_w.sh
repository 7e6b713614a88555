# round 2, GPU call W: flat plan builder with the warp-sorted runs + rank merge in place of the bitonic sort
cd /root/repo; mkdir -p gpurun_out
python -m pytest tests/test_gpu_tiles.py tests/test_gpu_hostio.py -q -m gpu -x 2>&1 | grep -v "^$" | tail -8 > gpurun_out/w_tests.log
python -m pytest tests/test_gpu_fullsize.py -q -m gpu -x -k "c5 or plj" 2>&1 | grep -v "^$" | tail -6 >> gpurun_out/w_tests.log
for w in dpd plj4m eydpd; do python tools/bench_nlist.py $w; done > gpurun_out/w_nlist.log 2>&1
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -k regex:"tiles_" -c 12 --csv --log-file gpurun_out/w_launches_eydpd.csv python tools/bench_nlist.py eydpd > gpurun_out/w_ncu1.log 2>&1
tail -n 6 gpurun_out/w_tests.log; cat gpurun_out/w_nlist.log; awk -F'","' '{print substr($5,1,50), $(NF-2), $NF}' gpurun_out/w_launches_eydpd.csv | tail -16
