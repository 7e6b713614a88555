cd /root/repo; mkdir -p gpurun_out
tr() { n=$1; shift; timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 10 --warmup 3 --no-cpu-baseline "$@" 2>> gpurun_out/r2b_err.log | tail -1 >> gpurun_out/r2b_bench.log; }
tr 8 --workload eydpd
tr 8 --workload eydpd --strong
tr 8 --workload dpd --strong --n 8000000
tr 8 --workload plj4m --strong
