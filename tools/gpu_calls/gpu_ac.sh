# round 2, GPU call AC: whole GPU suite on the head, default bench line, f64 A/B of the DPD stack kernel, ncu of the DPD stack kernel
cd /root/repo; mkdir -p gpurun_out
python -m pytest tests -q -m gpu -x 2>&1 | grep -v "^$" | tail -12 > gpurun_out/ac_tests.log
python bench.py > gpurun_out/ac_bench.json 2> gpurun_out/ac_bench.err; echo "rc=$?" >> gpurun_out/ac_bench.err
AZB_DPD_STACK=0 python bench.py --bits 64 --steps 5 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/ac_f64_ring.json 2> gpurun_out/ac_f64_ring.err
python bench.py --bits 64 --steps 5 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/ac_f64_stack.json 2> gpurun_out/ac_f64_stack.err
ncu --set full --clock-control none --import-source on -k regex:"dpd_stack_tile_kernel" --launch-skip 2 --launch-count 1 -o gpurun_out/ac_dpd_c5 -f python bench.py --steps 3 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/ac_ncu1.log 2>&1
python tools/ncu_summary.py gpurun_out/ac_dpd_c5.ncu-rep gpurun_out/ac_dpd_c5_summary.json "C5 8M f32: dpd_stack_tile_kernel (pruned view, one lane per particle)" > gpurun_out/ac_dpd_c5_summary.txt 2>&1
python tools/ncu_source.py gpurun_out/ac_dpd_c5.ncu-rep > gpurun_out/ac_dpd_c5_source.txt 2>&1
ncu --set full --clock-control none --import-source on -k regex:"dpd_stack_tile_kernel" --launch-skip 2 --launch-count 1 -o gpurun_out/ac_dpd_c2 -f python bench.py --workload dpd --steps 3 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/ac_ncu2.log 2>&1
python tools/ncu_summary.py gpurun_out/ac_dpd_c2.ncu-rep gpurun_out/ac_dpd_c2_summary.json "C2 1M f32: dpd_stack_tile_kernel (two lanes per particle)" > gpurun_out/ac_dpd_c2_summary.txt 2>&1
python tools/ncu_source.py gpurun_out/ac_dpd_c2.ncu-rep > gpurun_out/ac_dpd_c2_source.txt 2>&1
python tools/ncu_roofs.py gpurun_out/ac_dpd_c2_summary.json dpd_stack_tile_kernel DPDGeneralWeight gpurun_out/roofs_dpd_f32.json > /dev/null 2>&1
rm -f gpurun_out/ac_dpd_c5.ncu-rep gpurun_out/ac_dpd_c2.ncu-rep
tail -4 gpurun_out/ac_tests.log; tail -2 gpurun_out/ac_bench.err
python - <<'PY'
import json
for f in ("ac_bench","ac_f64_ring","ac_f64_stack"):
    d=json.loads([l for l in open(f"gpurun_out/{f}.json") if l.startswith('{')][-1])
    print(f, round(d["ms_per_step"],4), d["value"], [(k["kernel"],round(k["kernel_ms"],4),round(k["frac"],3)) for k in d["roofline"]["kernels"]], "e2e", d["e2e"]["ms_per_step"])
PY
