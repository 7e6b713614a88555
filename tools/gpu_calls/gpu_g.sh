# round 2, GPU call G: ncu captures of the C5 kernels (+ C2), summarised on the box (the .ncu-rep files are too large to bring
# back together); compute-sanitizer passes
cd /root/repo; mkdir -p gpurun_out
python bench.py --steps 3 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/g_pre.json 2> gpurun_out/g_pre.err || exit 1
ncu --set full --clock-control none --import-source on -k regex:"dpd_tile_kernel|pair_tile_kernel" --launch-skip 4 --launch-count 2 -o gpurun_out/g_c5 -f python bench.py --steps 3 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/g_ncu_c5.log 2>&1
python tools/ncu_summary.py gpurun_out/g_c5.ncu-rep gpurun_out/g_c5_summary.json "C5 8M f32: dpd_tile_kernel (pruned view) + pair_tile_kernel<ExpandedYukawa>" > gpurun_out/g_c5_summary.txt 2>&1
python tools/ncu_source.py gpurun_out/g_c5.ncu-rep > gpurun_out/g_c5_source_first_kernel.txt 2>&1
ncu --set full --clock-control none --import-source on -k regex:"dpd_tile_kernel" --launch-skip 2 --launch-count 1 -o gpurun_out/g_c2 -f python bench.py --workload dpd --steps 3 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/g_ncu_c2.log 2>&1
python tools/ncu_summary.py gpurun_out/g_c2.ncu-rep gpurun_out/g_c2_summary.json "C2 1M f32: dpd_tile_kernel" > gpurun_out/g_c2_summary.txt 2>&1
python tools/ncu_source.py gpurun_out/g_c2.ncu-rep > gpurun_out/g_c2_source.txt 2>&1
ncu --set full --clock-control none -k regex:"nlist_kernel|tiles_build|nlist_compact" --launch-skip 2 --launch-count 5 -o gpurun_out/g_rebuild -f python tools/bench_nlist.py eydpd > gpurun_out/g_ncu_rebuild.log 2>&1
python tools/ncu_summary.py gpurun_out/g_rebuild.ncu-rep gpurun_out/g_rebuild_summary.json "C5 8M rebuild kernels" > gpurun_out/g_rebuild_summary.txt 2>&1
rm -f gpurun_out/g_c5.ncu-rep gpurun_out/g_rebuild.ncu-rep
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 3 python tools/sanitizer_workload.py > gpurun_out/g_memcheck.log 2>&1; echo "memcheck rc=$?" >> gpurun_out/g_memcheck.log
timeout 900 compute-sanitizer --tool racecheck --error-exitcode 3 python tools/sanitizer_workload.py > gpurun_out/g_racecheck.log 2>&1; echo "racecheck rc=$?" >> gpurun_out/g_racecheck.log
tail -n 3 gpurun_out/g_memcheck.log; tail -n 3 gpurun_out/g_racecheck.log; ls -la gpurun_out/
