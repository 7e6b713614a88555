# round 2, GPU call C: whole GPU test suite (no -x), launch list of the default bench
cd /root/repo; mkdir -p gpurun_out
python -m pytest tests -q -m gpu -s 2>&1 | grep -v "^$" | tail -60 > gpurun_out/c_tests.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/c_launches.csv python bench.py --steps 3 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/c_ncu.log 2>&1
tail -5 gpurun_out/c_tests.log
