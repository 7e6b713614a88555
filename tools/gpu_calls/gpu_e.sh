# round 2, GPU call E: single-walk list build + pruned plans: tests, rebuild timings, bench
cd /root/repo; mkdir -p gpurun_out
python -m pytest tests -q -m gpu -x 2>&1 | grep -v "^$" | tail -30 > gpurun_out/e_tests.log
for w in dpd plj4m eydpd; do python tools/bench_nlist.py $w; done > gpurun_out/e_nlist.log 2>&1
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/e_bench.json 2> gpurun_out/e_bench.err
tail -6 gpurun_out/e_tests.log; cat gpurun_out/e_nlist.log; tail -3 gpurun_out/e_bench.err
