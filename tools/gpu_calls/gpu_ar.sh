# round 2, GPU call AR: tapered input chunks of the pipelined host step (e2e): equal vs tapered, chunk counts
cd /root/repo; mkdir -p gpurun_out
python -m pytest tests/test_gpu_hostio.py -q -m gpu -x 2>&1 | grep -v "^$" | tail -2
run() {
name=$1; shift
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras "$@" > gpurun_out/ar_$name.json 2> gpurun_out/ar_$name.err
python - $name <<'PY'
import json,sys
n=sys.argv[1]
try:
    d=json.loads([l for l in open(f"gpurun_out/ar_{n}.json") if l.startswith('{')][-1])
    print(n, "e2e ms", round(d["e2e"]["ms_per_step"],3), "G", round(d["e2e"]["value"]/1e9,3), "device ms", round(d["ms_per_step"],4))
except Exception as e: print(n, "failed", e, open(f"gpurun_out/ar_{n}.err").read()[-300:])
PY
}
run equal8 --e2e-taper none
run taper8
run equal8b --e2e-taper none
run taper8b
run taper10 --e2e-chunks 10 --e2e-taper 8,4,2,1
run taper12 --e2e-chunks 12 --e2e-taper 8,4,2,1
run taper8s --e2e-taper 6,3,1.5
run dpd_equal --workload dpd --e2e-taper none
run dpd_taper --workload dpd
