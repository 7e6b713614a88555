# round 2, GPU call P: ncu of the rebuild kernels (launch list + full capture with source of the list walk and the plan builder)
cd /root/repo; mkdir -p gpurun_out
python tools/bench_nlist.py eydpd > gpurun_out/p_pre.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/p_launches_eydpd.csv python tools/bench_nlist.py eydpd > gpurun_out/p_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"tiles_build_flat" --launch-skip 2 --launch-count 2 -o gpurun_out/p_tiles -f python tools/bench_nlist.py eydpd > gpurun_out/p_ncu2.log 2>&1
python tools/ncu_summary.py gpurun_out/p_tiles.ncu-rep gpurun_out/p_tiles_summary.json "C5 8M flat tile-plan builder" > gpurun_out/p_tiles_summary.txt 2>&1
python tools/ncu_source.py gpurun_out/p_tiles.ncu-rep > gpurun_out/p_tiles_source.txt 2>&1
ncu --set full --clock-control none --import-source on -k regex:"nlist_bcast" --launch-skip 2 --launch-count 1 -o gpurun_out/p_bcast -f python tools/bench_nlist.py eydpd > gpurun_out/p_ncu3.log 2>&1
python tools/ncu_summary.py gpurun_out/p_bcast.ncu-rep gpurun_out/p_bcast_summary.json "C5 8M broadcast list walk" > gpurun_out/p_bcast_summary.txt 2>&1
python tools/ncu_source.py gpurun_out/p_bcast.ncu-rep > gpurun_out/p_bcast_source.txt 2>&1
ls -la gpurun_out/
