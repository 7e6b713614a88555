# round 2, GPU call AO: DPD stack kernel with the slot-packet load kept one pass ahead: parity + timings
cd /root/repo; mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_golden_fixtures.py tests/test_gpu_tiles.py tests/test_gpu_hostio.py -q -m gpu -x 2>&1 | grep -v "^$" | tail -3 > gpurun_out/ao_tests.log
python -m pytest tests/test_gpu_fullsize.py -q -m gpu -x -s -k "c5 or dpd" 2>&1 | grep -v "^$" | tail -6 >> gpurun_out/ao_tests.log
cat gpurun_out/ao_tests.log
for w in eydpd dpd; do
python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/ao_$w.json 2> gpurun_out/ao_$w.err
python - $w <<'PY'
import json,sys
w=sys.argv[1]
d=json.loads([l for l in open(f"gpurun_out/ao_{w}.json") if l.startswith('{')][-1])
print(w, round(d["ms_per_step"],4), [(k["kernel"],round(k["kernel_ms"],4),round(k["frac"],3)) for k in d["roofline"]["kernels"]])
PY
done
