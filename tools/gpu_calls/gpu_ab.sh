# round 2, GPU call AB: DPD stack kernel v2 (8-byte slot loads, two pairs per pass) — timings
cd /root/repo; mkdir -p gpurun_out
for t in 1 2; do AZB_DPD_STACK=$t python -m pytest tests/test_gpu_parity.py tests/test_golden_fixtures.py -q -m gpu -x -k "dpd or DPD or c5 or C5 or eydpd" 2>&1 | grep -v "^$" | tail -2; done > gpurun_out/ab_tests.log 2>&1
for w in dpd eydpd; do
  for t in 1 2 4; do AZB_DPD_STACK=$t python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/ab_${w}_s$t.json 2> gpurun_out/ab_${w}_s$t.err; done
  AZB_DPD_STACK=1 python bench.py --workload $w --tile-size 256 --steps 10 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/ab_${w}_s1t256.json 2> gpurun_out/ab_${w}_s1t256.err
  AZB_DPD_STACK=2 python bench.py --workload $w --tile-size 64 --steps 10 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/ab_${w}_s2t64.json 2> gpurun_out/ab_${w}_s2t64.err
done
cat gpurun_out/ab_tests.log
python - <<'PY'
import json
for w in ("dpd","eydpd"):
    for s in ("s1","s2","s4","s1t256","s2t64"):
        try:
            d=json.loads([l for l in open(f"gpurun_out/ab_{w}_{s}.json") if l.startswith('{')][-1])
            print(w, s, round(d["ms_per_step"],4), [(k["kernel"],round(k["kernel_ms"],4),round(k["frac"],3)) for k in d["roofline"]["kernels"]])
        except Exception as e: print(w, s, "failed", e)
PY
