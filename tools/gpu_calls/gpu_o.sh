# round 2, GPU call O: masked broadcast list walk + aniso ring kernel — correctness, timings A/B
cd /root/repo; mkdir -p gpurun_out
python -m pytest tests/test_gpu_nlist.py tests/test_gpu_md.py tests/test_gpu_tiles.py tests/test_zz_no_alloc.py -q -m gpu -x 2>&1 | grep -v "^$" | tail -15 > gpurun_out/o_tests.log
python -m pytest tests/test_gpu_parity.py tests/test_golden_fixtures.py -q -m gpu -x 2>&1 | grep -v "^$" | tail -15 >> gpurun_out/o_tests.log
python -m pytest tests/test_gpu_fullsize.py -q -m gpu -x 2>&1 | grep -v "^$" | tail -6 >> gpurun_out/o_tests.log
for w in dpd plj4m eydpd; do python tools/bench_nlist.py $w; done > gpurun_out/o_nlist.log 2>&1
python bench.py --workload patchy --steps 20 --warmup 5 --no-cpu-baseline --no-extras > gpurun_out/o_patchy_ring.json 2> gpurun_out/o_patchy_ring.err
AZB_ANISO_NO_RING=1 python bench.py --workload patchy --steps 20 --warmup 5 --no-cpu-baseline --no-extras > gpurun_out/o_patchy_branch.json 2> gpurun_out/o_patchy_branch.err
for t in 2 8; do python bench.py --workload patchy --tpp $t --steps 20 --warmup 5 --no-cpu-baseline --no-extras > gpurun_out/o_patchy_ring_tpp$t.json 2> gpurun_out/o_patchy_ring_tpp$t.err; done
tail -n 12 gpurun_out/o_tests.log; cat gpurun_out/o_nlist.log
python - <<'PY'
import json
for f in ("o_patchy_ring","o_patchy_branch","o_patchy_ring_tpp2","o_patchy_ring_tpp8"):
    try:
        d=json.load(open(f"gpurun_out/{f}.json"))
        print(f, round(d["ms_per_step"],4), [(k["kernel"],round(k["kernel_ms"],4),round(k["frac"],3)) for k in d["roofline"]["kernels"]])
    except Exception as e: print(f, "failed", e)
PY
