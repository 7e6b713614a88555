#!/bin/bash
# scratch GPU script
cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/t.log
{
python bench.py --workload plj4m --tpp 2 --steps 20 --warmup 5 --no-cpu-baseline
AZB_NO_PACKED=1 python bench.py --workload plj4m --tpp 2 --steps 20 --warmup 5 --no-cpu-baseline
python bench.py --steps 30 --warmup 5 --no-cpu-baseline
python bench.py --workload patchy2m --steps 20 --warmup 5 --no-cpu-baseline
} > gpurun_out/b.log 2>&1
