# round 2, GPU call AM: where the interior-tile walk loses its gain (staging test vs loops): forced modes + a source-level capture
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
$B > gpurun_out/am_default.json 2> gpurun_out/am_default.err; show am_default
AZB_EXPERIMENT_QNEAR=-1 $B > gpurun_out/am_allwrap.json 2> gpurun_out/am_allwrap.err; show am_allwrap
AZB_EXPERIMENT_QNEAR=1e30 $B > gpurun_out/am_nowrap.json 2> gpurun_out/am_nowrap.err; show am_nowrap
ncu --set full --clock-control none --import-source on -k regex:"packed_pair_tile_kernel" --launch-skip 2 --launch-count 1 -o gpurun_out/am_ey -f python bench.py --steps 3 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/am_ncu1.log 2>&1
