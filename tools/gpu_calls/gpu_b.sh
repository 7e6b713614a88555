# round 2, GPU call B (2 GPUs): multi-GPU parity, MD with migration, bench at N=2
cd /root/repo; mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
echo skipped > gpurun_out/b_force.log
timeout 600 $TR --master-port 29534 tests/mgpu_md_check.py > gpurun_out/b_md.log 2>&1; echo "md rc=$?" >> gpurun_out/b_md.log
timeout 600 $TR --master-port 29535 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/b_bench2.json 2> gpurun_out/b_bench2.err; echo "rc=$?" >> gpurun_out/b_bench2.err
timeout 400 $TR --master-port 29536 bench.py --gpus 2 --workload melt --steps 5 --warmup 3 --no-extras > gpurun_out/b_melt2.json 2> gpurun_out/b_melt2.err; echo "rc=$?" >> gpurun_out/b_melt2.err
tail -4 gpurun_out/b_force.log; tail -6 gpurun_out/b_md.log; tail -3 gpurun_out/b_bench2.err; head -c 600 gpurun_out/b_bench2.json; tail -3 gpurun_out/b_melt2.err
