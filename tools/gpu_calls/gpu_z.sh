# round 2, GPU call Z: C1 1000-step trajectory parity (new test), e2e chunk counts
cd /root/repo; mkdir -p gpurun_out
python -m pytest tests/test_gpu_md.py -q -m gpu -x -s 2>&1 | grep -v "^$" | tail -20 > gpurun_out/z_tests.log
for c in 16 32; do python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras --e2e-chunks $c > gpurun_out/z_bench_c$c.json 2> gpurun_out/z_bench_c$c.err; done
cat gpurun_out/z_tests.log
python - <<'PY'
import json
for c in (16,32):
    d=json.loads([l for l in open(f"gpurun_out/z_bench_c{c}.json") if l.startswith('{')][-1])
    print(c, d["e2e"]["ms_per_step"], d["e2e"]["value"])
PY
