CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-extras --tpp 4"
$CMD > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:dpd_force_kernel -s 3 -c 2 -o gpurun_out/prof_dpd_q1 $CMD > gpurun_out/ncu.log 2>&1
