# round 2, GPU call AI (re-run after the ExpandedYukawa face test): final state — whole GPU suite, smoke, both bench arms as the driver runs them, launch list of the default bench
cd /root/repo; mkdir -p gpurun_out
( time python -m pytest tests -x -q -m gpu -s ) > gpurun_out/ai_tests.log 2>&1
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/ai_smoke.log 2>&1
( time python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 ) > gpurun_out/ai_ref.json 2> gpurun_out/ai_ref.err
( time python bench.py --gpus 1 --steps 20 --warmup 5 ) > gpurun_out/ai_bench.json 2> gpurun_out/ai_bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/ai_launches_c5.csv python bench.py --steps 3 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/ai_ncu0.log 2>&1
grep -E "passed|failed|fullsize parity" gpurun_out/ai_tests.log | tail -14; tail -2 gpurun_out/ai_smoke.log
python - <<'PY'
import json
d=json.loads([l for l in open("gpurun_out/ai_bench.json") if l.startswith('{')][-1])
print(round(d["ms_per_step"],4), d["value"], [(k["kernel"],round(k["kernel_ms"],4),round(k["frac"],3)) for k in d["roofline"]["kernels"]], "e2e", round(d["e2e"]["ms_per_step"],3), d["roofline"].get("binding_roof"))
for o in d.get("others",[]): print(o.get("workload","")[:50], o.get("ms_per_step"), [(k["kernel"],round(k["kernel_ms"],4),round(k["frac"],3)) for k in o.get("kernels",[])])
print("f64", d["f64"]["ms_per_step"], "amortised", d["amortised"]["ms_per_step"], d["amortised"]["md_step_ms"])
PY
tail -3 gpurun_out/ai_bench.err gpurun_out/ai_ref.err
ncu --set full --clock-control none --import-source on -k regex:"packed_pair_tile_kernel" --launch-skip 2 --launch-count 1 -o gpurun_out/ai_ey -f python bench.py --steps 3 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/ai_ncu1.log 2>&1
