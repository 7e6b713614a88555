# round 2, GPU call AS (2 GPUs): the tapered host step across the ghost exchange (bench at N = 2 checks e2e forces bit for bit)
cd /root/repo; mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 600 $TR --nproc-per-node 2 --master-port 29551 bench.py --gpus 2 --steps 20 --warmup 5 --no-extras > gpurun_out/as_bench2.json 2> gpurun_out/as_bench2.err; echo "rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open("gpurun_out/as_bench2.json") if l.startswith('{')][-1])
print(d["n_gpus"], d["value"], d["ms_per_step"], d["mgpu_check"]["ok"], d["e2e"]["value"], d["e2e"]["ms_per_step"])
PY
tail -3 gpurun_out/as_bench2.err
