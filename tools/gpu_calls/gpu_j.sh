cd /root/repo; mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_golden_fixtures.py tests/test_gpu_hostio.py -q -m gpu -x 2>&1 | grep -v "^$" | tail -12 > gpurun_out/j_tests.log
python -m pytest tests/test_gpu_fullsize.py -q -m gpu -x -k "c5" -s 2>&1 | grep -v "^$" | tail -6 >> gpurun_out/j_tests.log
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/j_bench.json 2> gpurun_out/j_bench.err
AZB_NO_PACKED=1 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/j_bench_generic.json 2> gpurun_out/j_bench_generic.err
for c in 16 32; do python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras --e2e-chunks $c > gpurun_out/j_bench_c$c.json 2> gpurun_out/j_bench_c$c.err; done
tail -n 8 gpurun_out/j_tests.log
python - <<'PY'
import json
for f in ("j_bench","j_bench_generic","j_bench_c16","j_bench_c32"):
    try:
        d=json.load(open(f"gpurun_out/{f}.json"))
        print(f, round(d["ms_per_step"],4), [(k["kernel"],round(k["kernel_ms"],4),round(k["frac"],3)) for k in d["roofline"]["kernels"]], "e2e", round(d["e2e"]["ms_per_step"],3))
    except Exception as e: print(f, "failed", e)
PY
