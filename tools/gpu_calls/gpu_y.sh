# round 2, GPU call Y: slot rows in ascending slot order (sort_rows) — correctness and kernel / rebuild timings A/B
cd /root/repo; mkdir -p gpurun_out
python -m pytest tests/test_gpu_tiles.py -q -m gpu -x 2>&1 | grep -v "^$" | tail -8 > gpurun_out/y_tests.log
AZB_SORT_ROWS=1 python -m pytest tests/test_gpu_parity.py tests/test_golden_fixtures.py tests/test_gpu_hostio.py tests/test_gpu_md.py -q -m gpu -x 2>&1 | grep -v "^$" | tail -8 >> gpurun_out/y_tests.log
AZB_SORT_ROWS=1 python -m pytest tests/test_gpu_fullsize.py -q -m gpu -x 2>&1 | grep -v "^$" | tail -6 >> gpurun_out/y_tests.log
for w in eydpd plj4m dpd patchy melt colloid hertz; do
  for s in 0 1; do AZB_SORT_ROWS=$s python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/y_${w}_s$s.json 2> gpurun_out/y_${w}_s$s.err; done
done
for s in 0 1; do AZB_SORT_ROWS=$s python tools/bench_nlist.py eydpd > gpurun_out/y_nlist_s$s.log 2>&1; done
tail -n 10 gpurun_out/y_tests.log; cat gpurun_out/y_nlist_s0.log gpurun_out/y_nlist_s1.log
python - <<'PY'
import json
for w in "eydpd plj4m dpd patchy melt colloid hertz".split():
    for s in (0,1):
        try:
            d=json.loads([l for l in open(f"gpurun_out/y_{w}_s{s}.json") if l.startswith('{')][-1])
            print(w, s, round(d["ms_per_step"],4), [(k["kernel"],round(k["kernel_ms"],4),round(k["frac"],3)) for k in d["roofline"]["kernels"]], "e2e", round(d["e2e"]["ms_per_step"],3), d["neighbor_data"][60:140])
        except Exception as e: print(w, s, "failed", e)
PY
