# round 2, GPU call K (8 GPUs): multi-GPU parity + MD with migration at 8 and 4 ranks, bench at N = 8 and 4 (weak + strong record)
cd /root/repo; mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 500 $TR --nproc-per-node 8 --master-port 29533 tests/mgpu_force_check.py > gpurun_out/k_force8.log 2>&1; echo "force8 rc=$?" >> gpurun_out/k_force8.log
timeout 700 $TR --nproc-per-node 8 --master-port 29534 tests/mgpu_md_check.py > gpurun_out/k_md8.log 2>&1; echo "md8 rc=$?" >> gpurun_out/k_md8.log
timeout 700 $TR --nproc-per-node 4 --master-port 29537 tests/mgpu_md_check.py > gpurun_out/k_md4.log 2>&1; echo "md4 rc=$?" >> gpurun_out/k_md4.log
timeout 800 $TR --nproc-per-node 8 --master-port 29535 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/k_bench8.json 2> gpurun_out/k_bench8.err; echo "rc=$?" >> gpurun_out/k_bench8.err
timeout 800 $TR --nproc-per-node 4 --master-port 29536 bench.py --gpus 4 --steps 20 --warmup 5 > gpurun_out/k_bench4.json 2> gpurun_out/k_bench4.err; echo "rc=$?" >> gpurun_out/k_bench4.err
timeout 500 $TR --nproc-per-node 8 --master-port 29538 bench.py --gpus 8 --workload melt --steps 5 --warmup 3 --no-extras > gpurun_out/k_melt8.json 2> gpurun_out/k_melt8.err; echo "rc=$?" >> gpurun_out/k_melt8.err
grep "ok\|rc=" gpurun_out/k_force8.log gpurun_out/k_md8.log gpurun_out/k_md4.log | tail -12; tail -n 2 gpurun_out/k_bench8.err gpurun_out/k_bench4.err gpurun_out/k_melt8.err
