# round 2, GPU call K (8 GPUs): multi-GPU parity + MD with migration at 8 ranks, bench at N = 8 (weak + strong record, e2e, mgpu_check)
cd /root/repo; mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 $TR --nproc-per-node 8 --master-port 29533 tests/mgpu_force_check.py > gpurun_out/k_force8.log 2>&1; echo "force8 rc=$?" >> gpurun_out/k_force8.log
timeout 400 $TR --nproc-per-node 8 --master-port 29534 tests/mgpu_md_check.py > gpurun_out/k_md8.log 2>&1; echo "md8 rc=$?" >> gpurun_out/k_md8.log
timeout 600 $TR --nproc-per-node 8 --master-port 29535 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/k_bench8.json 2> gpurun_out/k_bench8.err; echo "rc=$?" >> gpurun_out/k_bench8.err
grep "ok\|rc=" gpurun_out/k_force8.log gpurun_out/k_md8.log | tail -8; tail -n 2 gpurun_out/k_bench8.err; grep "^{" gpurun_out/k_bench8.json | head -c 1200
