# round 2, GPU call A: whole GPU test suite, default bench (both arms), launch list
cd /root/repo; mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu -s 2>&1 | grep -v "^$" | tail -40 > gpurun_out/a_tests.log
python bench.py --steps 20 --warmup 5 > gpurun_out/a_bench.json 2> gpurun_out/a_bench.err
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/a_ref.json 2> gpurun_out/a_ref.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/a_launches.csv python bench.py --steps 3 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/a_ncu.log 2>&1
tail -3 gpurun_out/a_tests.log; head -c 1500 gpurun_out/a_bench.json; tail -3 gpurun_out/a_bench.err
