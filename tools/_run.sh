python bench.py --steps 20 --warmup 5 > gpurun_out/r1l_bench_dpd.json 2> gpurun_out/r1l_err.log
python bench.py --steps 20 --warmup 5 --e2e-chunks 4 --no-cpu-baseline > gpurun_out/r1l_bench_dpd_c4.json 2>> gpurun_out/r1l_err.log
python bench.py --workload plj4m --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r1l_bench_plj4m.json 2>> gpurun_out/r1l_err.log
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r1l_bench_ref.json 2>> gpurun_out/r1l_err.log
nvidia-smi topo -m > gpurun_out/r1l_topo.txt 2>&1
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-extras --e2e-chunks 1"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1l_launches.csv $CMD > gpurun_out/ncu1.log 2>&1
$CMD > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:dpd_tile_kernel -s 3 -c 2 -o gpurun_out/prof_dpd_r1l $CMD > gpurun_out/ncu2.log 2>&1
