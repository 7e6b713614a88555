cd /root/repo; mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -6) > gpurun_out/s14j_tests.log
timeout 300 python bench.py > gpurun_out/s14j_bench_dpd.json 2> gpurun_out/s14j_bench_dpd.err
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/s14j_bench_ref.json 2> gpurun_out/s14j_bench_ref.err
timeout 300 python bench.py --workload plj4m --cpu-seconds 5 > gpurun_out/s14j_bench_plj4m.json 2> gpurun_out/s14j_bench_plj4m.err
timeout 300 python bench.py --workload colloid --cpu-seconds 5 > gpurun_out/s14j_bench_colloid.json 2> gpurun_out/s14j_bench_colloid.err
# launch list of the default bench command (cold-cache, serialised: shares, not absolutes)
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/s14j_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/s14j_ncu_list.log 2>&1
# full capture of the LJ kernel at the tile size the force now picks
ncu --set full --clock-control none --import-source on -k regex:plj_packed -c 1 -s 4 -o gpurun_out/s14j_plj256 -f python bench.py --workload plj4m --steps 3 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/s14j_ncu_plj.log 2>&1
cat gpurun_out/s14j_tests.log; cut -c1-300 gpurun_out/s14j_bench_dpd.json; cut -c1-600 gpurun_out/s14j_bench_ref.json; tail -2 gpurun_out/s14j_ncu_plj.log
