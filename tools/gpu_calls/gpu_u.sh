# round 2, GPU call U: ncu of the plan kernels (launch list + full capture with source)
cd /root/repo; mkdir -p gpurun_out
ncu --metrics gpu__time_duration.sum --clock-control none -c 160 --csv --log-file gpurun_out/u_launches_eydpd.csv python tools/bench_nlist.py eydpd > gpurun_out/u_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"tiles_prune" --launch-skip 1 --launch-count 1 -o gpurun_out/u_prune -f python tools/bench_nlist.py eydpd > gpurun_out/u_ncu2.log 2>&1
python tools/ncu_summary.py gpurun_out/u_prune.ncu-rep gpurun_out/u_prune_summary.json "C5 8M derived pruned plan" > gpurun_out/u_prune_summary.txt 2>&1
python tools/ncu_source.py gpurun_out/u_prune.ncu-rep > gpurun_out/u_prune_source.txt 2>&1
ncu --set full --clock-control none --import-source on -k regex:"tiles_build_flat" --launch-skip 1 --launch-count 1 -o gpurun_out/u_flat -f python tools/bench_nlist.py eydpd > gpurun_out/u_ncu3.log 2>&1
python tools/ncu_summary.py gpurun_out/u_flat.ncu-rep gpurun_out/u_flat_summary.json "C5 8M flat plan builder (no pruning)" > gpurun_out/u_flat_summary.txt 2>&1
python tools/ncu_source.py gpurun_out/u_flat.ncu-rep > gpurun_out/u_flat_source.txt 2>&1
rm -f gpurun_out/u_flat.ncu-rep gpurun_out/u_prune.ncu-rep
ls -la gpurun_out/ | tail -12
