# round 2, GPU call AE: stack kernel with hoisted header loads — parity + timings
cd /root/repo; mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_golden_fixtures.py tests/test_gpu_tiles.py tests/test_gpu_hostio.py -q -m gpu -x 2>&1 | grep -v "^$" | tail -3 > gpurun_out/ae_tests.log
python -m pytest tests/test_gpu_fullsize.py -q -m gpu -x -k "dpd or c5" 2>&1 | grep -v "^$" | tail -3 >> gpurun_out/ae_tests.log
for w in dpd eydpd; do
  python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/ae_${w}.json 2> gpurun_out/ae_${w}.err
  python bench.py --workload $w --bits 64 --steps 5 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/ae_${w}_f64.json 2> gpurun_out/ae_${w}_f64.err
done
cat gpurun_out/ae_tests.log
python - <<'PY'
import json
for f in ("ae_dpd","ae_dpd_f64","ae_eydpd","ae_eydpd_f64"):
    d=json.loads([l for l in open(f"gpurun_out/{f}.json") if l.startswith('{')][-1])
    print(f, round(d["ms_per_step"],4), [(k["kernel"],round(k["kernel_ms"],4),round(k["frac"],3)) for k in d["roofline"]["kernels"]], "e2e", round(d["e2e"]["ms_per_step"],3))
PY
