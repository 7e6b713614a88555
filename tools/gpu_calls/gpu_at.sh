# round 2, GPU call AT: what bounds the C5 host step — PCIe probe of this box, chunk counts with equal chunks, issue trace
cd /root/repo; mkdir -p gpurun_out
python tools/pcie_probe.py 2>&1 | tail -4
run() {
name=$1; shift
AZB_TRACE_HOST_STEP=1 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras "$@" > gpurun_out/at_$name.json 2> gpurun_out/at_$name.err
python - $name <<'PY'
import json,sys,re
n=sys.argv[1]
try:
    d=json.loads([l for l in open(f"gpurun_out/at_{n}.json") if l.startswith('{')][-1])
    tr=[l for l in open(f"gpurun_out/at_{n}.err") if "azb_host_step" in l]
    print(n, "e2e ms", round(d["e2e"]["ms_per_step"],3), "G", round(d["e2e"]["value"]/1e9,3), tr[-1].strip() if tr else "")
except Exception as e: print(n, "failed", e, open(f"gpurun_out/at_{n}.err").read()[-300:])
PY
}
run k4 --e2e-chunks 4 --e2e-taper none
run k8 --e2e-chunks 8 --e2e-taper none
run k16 --e2e-chunks 16 --e2e-taper none
run k32 --e2e-chunks 32 --e2e-taper none
run k16t --e2e-chunks 16 --e2e-taper 4,2,1
