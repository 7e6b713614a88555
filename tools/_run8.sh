cd /root/repo; mkdir -p gpurun_out
tr() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $2 bench.py --gpus $1 --steps 20 --warmup 5 --no-cpu-baseline --no-extras "${@:3}" 2>> gpurun_out/s14u_err.log | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('N=$1 ${*:3}', d['ms_per_step'], d['value'], d['scaling'], 'e2e', d['e2e']['value'])" >> gpurun_out/s14u_scale.log 2>&1; }
python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-extras 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('N=1', d['ms_per_step'], d['value'])" >> gpurun_out/s14u_scale.log
tr 2 29601
tr 4 29602
tr 8 29603
tr 8 29604 --workload plj4m
tr 8 29605 --workload eydpd
tr 8 29606 --workload eydpd --strong
(timeout 300 python -m pytest tests/test_gpu_multi.py -x -q -m gpu 2>&1 | tail -3) >> gpurun_out/s14u_scale.log
cat gpurun_out/s14u_scale.log; tail -3 gpurun_out/s14u_err.log
