# round 2, GPU call AF: whole GPU suite at HEAD, smoke, both bench arms with wall time, MUFU micro-benchmark, ncu of the bench default
cd /root/repo; mkdir -p gpurun_out
( time python -m pytest tests -x -q -m gpu ) > gpurun_out/af_tests.log 2>&1
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/af_smoke.log 2>&1
tools/microbench/mufu > gpurun_out/af_mufu.txt 2>&1
( time python bench.py --impl reference --steps 3 --warmup 1 ) > gpurun_out/af_ref.json 2> gpurun_out/af_ref.err
( time python bench.py ) > gpurun_out/af_bench.json 2> gpurun_out/af_bench.err
AZB_NO_PACKED=1 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/af_nopacked.json 2> gpurun_out/af_nopacked.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/af_launches_c5.csv python bench.py --steps 3 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/af_ncu0.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"packed_pair_tile_kernel" --launch-skip 2 --launch-count 1 -o gpurun_out/af_ey -f python bench.py --steps 3 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/af_ncu1.log 2>&1
tail -3 gpurun_out/af_tests.log; tail -2 gpurun_out/af_smoke.log; cat gpurun_out/af_mufu.txt
python - <<'PY'
import json
for f in ("af_bench","af_nopacked"):
    try:
        d=json.loads([l for l in open(f"gpurun_out/{f}.json") if l.startswith('{')][-1])
        print(f, round(d["ms_per_step"],4), [(k["kernel"],round(k["kernel_ms"],4),round(k["frac"],3)) for k in d["roofline"]["kernels"]], "e2e", round(d["e2e"]["ms_per_step"],3) if "e2e" in d else None)
    except Exception as e: print(f, "failed", e)
PY
tail -4 gpurun_out/af_bench.err gpurun_out/af_ref.err
