# round 2, GPU call AK: upper bound of a walk without the per-visit wrap test (timing only: AZB_EXPERIMENT_NOWRAP is wrong on boundary tiles)
cd /root/repo; mkdir -p gpurun_out
show() {
python - "$1" <<'PY'
import json,sys
f=sys.argv[1]
try:
    d=json.loads([l for l in open(f"gpurun_out/{f}.json") if l.startswith('{')][-1])
    print(f, round(d["ms_per_step"],4), [(k["kernel"],round(k["kernel_ms"],4),round(k["frac"],3)) for k in d["roofline"]["kernels"]])
except Exception as e: print(f,"failed",e, open(f"gpurun_out/{f}.err").read()[-400:])
PY
}
B="python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras"
$B > gpurun_out/ak_base.json 2> gpurun_out/ak_base.err; show ak_base
AZB_EXPERIMENT_NOWRAP=1 $B > gpurun_out/ak_nowrap.json 2> gpurun_out/ak_nowrap.err; show ak_nowrap
$B > gpurun_out/ak_base2.json 2> gpurun_out/ak_base2.err; show ak_base2
AZB_EXPERIMENT_NOWRAP=1 $B > gpurun_out/ak_nowrap2.json 2> gpurun_out/ak_nowrap2.err; show ak_nowrap2
