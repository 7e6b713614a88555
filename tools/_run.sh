cd /root/repo; mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -30 > gpurun_out/r1t_pytest.log
for w in dpd plj4m; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --workload $w --steps 10 --warmup 3 --no-cpu-baseline 2>> gpurun_out/r1t_err.log | tail -1 >> gpurun_out/r1t_bench.log
done

