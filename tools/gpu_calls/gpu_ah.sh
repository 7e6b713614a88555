# round 2, GPU call AH: what moves the ExpandedYukawa kernel now that the L1 data pipe binds it (92 %): lanes per particle, sorted
# / bank-rotated slot rows, tile size
cd /root/repo; mkdir -p gpurun_out
run() { # name, env..., -- args
  name=$1; shift
  env "$@" > /dev/null 2>&1
}
show() {
python - "$1" <<'PY'
import json,sys
f=sys.argv[1]
try:
    d=json.loads([l for l in open(f"gpurun_out/{f}.json") if l.startswith('{')][-1])
    print(f, round(d["ms_per_step"],4), [(k["kernel"],round(k["kernel_ms"],4),round(k["frac"],3)) for k in d["roofline"]["kernels"]], d.get("neighbor_data","")[:120])
except Exception as e: print(f,"failed",e, open(f"gpurun_out/{f}.err").read()[-400:])
PY
}
B="python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras"
$B > gpurun_out/ah_base.json 2> gpurun_out/ah_base.err; show ah_base
$B --tpp 4 > gpurun_out/ah_tpp4.json 2> gpurun_out/ah_tpp4.err; show ah_tpp4
$B --tpp 2 > gpurun_out/ah_tpp2.json 2> gpurun_out/ah_tpp2.err; show ah_tpp2
AZB_SORT_ROWS=1 $B > gpurun_out/ah_sort.json 2> gpurun_out/ah_sort.err; show ah_sort
AZB_SORT_ROWS=1 $B --tpp 4 > gpurun_out/ah_sort_tpp4.json 2> gpurun_out/ah_sort_tpp4.err; show ah_sort_tpp4
AZB_BANK_ROTATE=1 $B > gpurun_out/ah_rot.json 2> gpurun_out/ah_rot.err; show ah_rot
$B --tile-size 64 > gpurun_out/ah_t64.json 2> gpurun_out/ah_t64.err; show ah_t64
$B --tile-size 256 > gpurun_out/ah_t256.json 2> gpurun_out/ah_t256.err; show ah_t256
$B --tile-size 256 --tpp 4 > gpurun_out/ah_t256_tpp4.json 2> gpurun_out/ah_t256_tpp4.err; show ah_t256_tpp4
