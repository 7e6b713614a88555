# round 2, GPU call AA: DPD tile kernel with lane-private stacks — parity and timings against the ring walk
cd /root/repo; mkdir -p gpurun_out
for t in 1 2 4; do AZB_DPD_STACK=$t python -m pytest tests/test_gpu_parity.py tests/test_golden_fixtures.py -q -m gpu -x -k "dpd or DPD or c5 or C5 or eydpd" 2>&1 | grep -v "^$" | tail -4; done > gpurun_out/aa_tests.log 2>&1
AZB_DPD_STACK=1 python -m pytest tests/test_gpu_fullsize.py tests/test_gpu_tiles.py -q -m gpu -x -k "dpd or c5 or pruned" 2>&1 | grep -v "^$" | tail -4 >> gpurun_out/aa_tests.log
for w in dpd eydpd; do
  python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/aa_${w}_ring.json 2> gpurun_out/aa_${w}_ring.err
  for t in 1 2 4; do AZB_DPD_STACK=$t python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/aa_${w}_s$t.json 2> gpurun_out/aa_${w}_s$t.err; done
done
cat gpurun_out/aa_tests.log
python - <<'PY'
import json
for w in ("dpd","eydpd"):
    for s in ("ring","s1","s2","s4"):
        try:
            d=json.loads([l for l in open(f"gpurun_out/aa_{w}_{s}.json") if l.startswith('{')][-1])
            print(w, s, round(d["ms_per_step"],4), [(k["kernel"],round(k["kernel_ms"],4),round(k["frac"],3)) for k in d["roofline"]["kernels"]])
        except Exception as e: print(w, s, "failed", e)
PY
