cd /root/repo; mkdir -p gpurun_out
run() { python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-extras "$@" 2>> gpurun_out/s14i_err.log | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$TAG $*', d['ms_per_step'], d['roofline']['kernel_ms'], d['value'], d['roofline']['frac'])" >> gpurun_out/s14i_sweep.log 2>&1; }
cp azplugins_b200/libazb200.so /tmp/lib_default.so
for v in default q2r64 q4m5 default; do
  if [ $v = default ]; then cp /tmp/lib_default.so azplugins_b200/libazb200.so; else cp azplugins_b200/libazb200_$v.so azplugins_b200/libazb200.so; fi
  TAG=$v run --workload dpd
done
cp /tmp/lib_default.so azplugins_b200/libazb200.so
cat gpurun_out/s14i_sweep.log; tail -5 gpurun_out/s14i_err.log
