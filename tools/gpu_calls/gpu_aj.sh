# round 2, GPU call AJ (2 GPUs): final state — multi-GPU tests, force and MD checks, both bench arms at N = 2
cd /root/repo; mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 python -m pytest tests/test_gpu_multi.py -q -m gpu -x 2>&1 | grep -v "^$" | tail -5 > gpurun_out/aj_multi_tests.log
timeout 400 $TR --nproc-per-node 2 --master-port 29533 tests/mgpu_force_check.py > gpurun_out/aj_force2.log 2>&1; echo "force2 rc=$?" >> gpurun_out/aj_force2.log
timeout 600 $TR --nproc-per-node 2 --master-port 29534 tests/mgpu_md_check.py > gpurun_out/aj_md2.log 2>&1; echo "md2 rc=$?" >> gpurun_out/aj_md2.log
timeout 800 $TR --nproc-per-node 2 --master-port 29535 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/aj_bench2.json 2> gpurun_out/aj_bench2.err; echo "rc=$?" >> gpurun_out/aj_bench2.err
tail -3 gpurun_out/aj_multi_tests.log; grep "ok\|rc=\|FAIL\|Error" gpurun_out/aj_force2.log gpurun_out/aj_md2.log | tail -12; tail -n 2 gpurun_out/aj_bench2.err; head -c 1500 gpurun_out/aj_bench2.json
timeout 300 $TR --nproc-per-node 2 --master-port 29536 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 > gpurun_out/aj_ref2.json 2> gpurun_out/aj_ref2.err; echo "ref rc=$?"; head -c 300 gpurun_out/aj_ref2.json
