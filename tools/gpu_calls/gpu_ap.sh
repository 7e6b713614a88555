# round 2, GPU call AP (4 GPUs): weak + strong record of the default bench with the round's last kernels
cd /root/repo; mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 600 $TR --nproc-per-node 4 --master-port 29541 bench.py --gpus 4 --steps 20 --warmup 5 > gpurun_out/ap_bench4.json 2> gpurun_out/ap_bench4.err; echo "rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open("gpurun_out/ap_bench4.json") if l.startswith('{')][-1])
print(d["n_gpus"], d["value"], d["ms_per_step"], d["mgpu_check"], d["e2e"]["value"], d.get("strong"))
PY
