# round 2, GPU call F: plans emitted by the list build
cd /root/repo; mkdir -p gpurun_out
python -m pytest tests/test_gpu_tiles.py tests/test_gpu_nlist.py -q -m gpu -x 2>&1 | grep -v "^$" | tail -30 > gpurun_out/f_tests.log
python -m pytest tests -q -m gpu -x 2>&1 | grep -v "^$" | tail -15 >> gpurun_out/f_tests.log
for w in dpd plj4m eydpd; do python tools/bench_nlist.py $w; done > gpurun_out/f_nlist.log 2>&1
AZB_FUSED_PRUNED=0 python tools/bench_nlist.py eydpd >> gpurun_out/f_nlist.log 2>&1
AZB_NO_FUSED_PLANS=1 python tools/bench_nlist.py eydpd >> gpurun_out/f_nlist.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/f_launches.csv python tools/bench_nlist.py eydpd > gpurun_out/f_ncu.log 2>&1
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/f_bench.json 2> gpurun_out/f_bench.err
AZB_FUSED_PRUNED=0 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/f_bench_ab.json 2> gpurun_out/f_bench_ab.err
tail -12 gpurun_out/f_tests.log; cat gpurun_out/f_nlist.log; tail -3 gpurun_out/f_bench.err
