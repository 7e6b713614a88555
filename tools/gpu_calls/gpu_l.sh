# round 2, GPU call L: whole GPU suite and the default bench line on the current head
cd /root/repo; mkdir -p gpurun_out
python -m pytest tests -q -m gpu -x 2>&1 | grep -v "^$" | tail -15 > gpurun_out/l_tests.log
python bench.py > gpurun_out/l_bench.json 2> gpurun_out/l_bench.err; echo "rc=$?" >> gpurun_out/l_bench.err
tail -6 gpurun_out/l_tests.log; tail -3 gpurun_out/l_bench.err; head -c 3000 gpurun_out/l_bench.json
