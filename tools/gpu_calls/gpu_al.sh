# round 2, GPU call AL: interior-tile walk without the wrap test (ExpandedYukawa packed kernel): parity + timing
cd /root/repo; mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_golden_fixtures.py tests/test_gpu_tiles.py tests/test_gpu_hostio.py -q -m gpu -x 2>&1 | grep -v "^$" | tail -3 > gpurun_out/al_tests.log
python -m pytest tests/test_gpu_fullsize.py -q -m gpu -x -s -k "c5" 2>&1 | grep -v "^$" | tail -6 >> gpurun_out/al_tests.log
cat gpurun_out/al_tests.log
python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-extras > gpurun_out/al_eydpd.json 2> gpurun_out/al_eydpd.err
python - <<'PY'
import json
d=json.loads([l for l in open("gpurun_out/al_eydpd.json") if l.startswith('{')][-1])
print(round(d["ms_per_step"],4), [(k["kernel"],round(k["kernel_ms"],4),round(k["frac"],3)) for k in d["roofline"]["kernels"]], "e2e", round(d["e2e"]["ms_per_step"],3))
PY
