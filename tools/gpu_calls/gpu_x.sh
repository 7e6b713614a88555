# round 2, GPU call X: ncu of the default bench (launch list, full capture of the two force kernels, roofs file), sanitizer passes
cd /root/repo; mkdir -p gpurun_out
python bench.py --steps 3 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/x_pre.json 2> gpurun_out/x_pre.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/x_launches_c5.csv python bench.py --steps 3 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/x_ncu0.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"packed_pair_tile_kernel|dpd_tile_kernel" --launch-skip 4 --launch-count 2 -o gpurun_out/x_c5 -f python bench.py --steps 3 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/x_ncu1.log 2>&1
python tools/ncu_summary.py gpurun_out/x_c5.ncu-rep gpurun_out/x_c5_summary.json "C5 8M f32: dpd_tile_kernel (pruned view) + packed_pair_tile_kernel<ExpandedYukawa>" > gpurun_out/x_c5_summary.txt 2>&1
python tools/ncu_roofs.py gpurun_out/x_c5_summary.json packed_pair_tile_kernel ExpandedYukawa gpurun_out/roofs_eydpd_f32.json > gpurun_out/x_roofs.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"packed_pair_tile_kernel" --launch-skip 2 --launch-count 1 -o gpurun_out/x_ey -f python bench.py --steps 3 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/x_ncu2.log 2>&1
python tools/ncu_source.py gpurun_out/x_ey.ncu-rep > gpurun_out/x_ey_source.txt 2>&1
rm -f gpurun_out/x_c5.ncu-rep gpurun_out/x_ey.ncu-rep
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 3 python tools/sanitizer_workload.py > gpurun_out/x_memcheck.log 2>&1; echo "memcheck rc=$?" >> gpurun_out/x_memcheck.log
timeout 900 compute-sanitizer --tool racecheck --error-exitcode 3 python tools/sanitizer_workload.py > gpurun_out/x_racecheck.log 2>&1; echo "racecheck rc=$?" >> gpurun_out/x_racecheck.log
tail -n 3 gpurun_out/x_memcheck.log; tail -n 3 gpurun_out/x_racecheck.log; cat gpurun_out/x_roofs.log | head -30
