# round 2, GPU call AG: ExpandedYukawa packed kernel with fewer FP32 operations per entry — parity, timings, capture
cd /root/repo; mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_golden_fixtures.py tests/test_gpu_tiles.py tests/test_gpu_hostio.py -q -m gpu -x 2>&1 | grep -v "^$" | tail -3 > gpurun_out/ag_tests.log
python -m pytest tests/test_gpu_fullsize.py -q -m gpu -x -s -k "c5" 2>&1 | grep -v "^$" | tail -6 >> gpurun_out/ag_tests.log
python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-extras > gpurun_out/ag_eydpd.json 2> gpurun_out/ag_eydpd.err
ncu --set full --clock-control none --import-source on -k regex:"packed_pair_tile_kernel" --launch-skip 2 --launch-count 1 -o gpurun_out/ag_ey -f python bench.py --steps 3 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/ag_ncu1.log 2>&1
ncu -i gpurun_out/ag_ey.ncu-rep --page raw --csv > gpurun_out/ag_ey_raw.csv 2>/dev/null
cat gpurun_out/ag_tests.log
python - <<'PY'
import json
for f in ("ag_eydpd",):
    d=json.loads([l for l in open(f"gpurun_out/{f}.json") if l.startswith('{')][-1])
    print(f, round(d["ms_per_step"],4), [(k["kernel"],round(k["kernel_ms"],4),round(k["frac"],3)) for k in d["roofline"]["kernels"]], "e2e", round(d["e2e"]["ms_per_step"],3))
PY
for s in default 0 2; do
  if [ $s = default ]; then unset AZB_DPD_STACK; else export AZB_DPD_STACK=$s; fi
  python bench.py --workload dpd --steps 10 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/ag_dpd_$s.json 2> gpurun_out/ag_dpd_$s.err
  python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/ag_dpd_$s.json") if l.startswith('{')][-1])
print("dpd stack=$s", round(d["ms_per_step"],4), [(k["kernel"],round(k["kernel_ms"],4),round(k["frac"],3)) for k in d["roofline"]["kernels"]])
PY
done
unset AZB_DPD_STACK
