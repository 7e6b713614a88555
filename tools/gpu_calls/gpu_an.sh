# round 2, GPU call AN: rows away from the periodic faces walk without the wrap test (both packed kernels): parity + timings
cd /root/repo; mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_golden_fixtures.py tests/test_gpu_tiles.py tests/test_gpu_hostio.py tests/test_gpu_md.py -q -m gpu -x 2>&1 | grep -v "^$" | tail -3 > gpurun_out/an_tests.log
python -m pytest tests/test_gpu_fullsize.py -q -m gpu -x -s -k "c5 or plj or c3" 2>&1 | grep -v "^$" | tail -9 >> gpurun_out/an_tests.log
cat gpurun_out/an_tests.log
for w in eydpd plj4m melt; do
python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/an_$w.json 2> gpurun_out/an_$w.err
python - $w <<'PY'
import json,sys
w=sys.argv[1]
d=json.loads([l for l in open(f"gpurun_out/an_{w}.json") if l.startswith('{')][-1])
print(w, round(d["ms_per_step"],4), [(k["kernel"],round(k["kernel_ms"],4),round(k["frac"],3)) for k in d["roofline"]["kernels"]])
PY
done
