"""Where the end-to-end step's time goes: azb_host_step with legs switched off (DPD C2 workload)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import azplugins_b200 as az
from azplugins_b200 import nlist, pair, workloads as W, _abi
from azplugins_b200.hostio import HostStep
import ctypes as C

snap = W.dpd_fluid()
nl = nlist.Cell(buffer=0.4)
f = pair.DPDGeneralWeight(nlist=nl, kT=1.0, default_r_cut=1.0)
f.dt = 0.01
for k, v in W.DPD_PARAMS.items():
    f.params[k] = v
st = az.State(snap, device="cuda", scalar_bits=32, seed=42)
f._attach(st)
f.compute(0)
torch.cuda.synchronize()


def timed(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e3


for chunks in (8, 16):
    hs = HostStep(f, chunks=chunks)
    h_in, h_out = hs.pinned_inputs(), hs.pinned_outputs()
    hs.run(0, h_in, h_out)
    a, whole, keep, outs, sched = hs._schedule
    ch = keep[1]
    full = timed(lambda: hs.run(1, h_in, h_out))
    waits = [s[0] for s in sched]
    n_out = a.n_outputs
    a.n_outputs = 0
    no_d2h = timed(lambda: hs.run(1, h_in, h_out))
    saved = [c.tile_count for c in ch]
    for c in ch:
        c.tile_count = 0
    only_h2d = timed(lambda: hs.run(1, h_in, h_out))
    for c, n in zip(ch, saved):
        c.tile_count = n
    a.n_outputs = n_out
    a.n_inputs = 0
    no_h2d = timed(lambda: hs.run(1, h_in, h_out))
    a.n_outputs = 0
    only_k = timed(lambda: hs.run(1, h_in, h_out))
    print(f"chunks={chunks}: full {full:.3f} ms | no D2H {no_d2h:.3f} | H2D only {only_h2d:.3f} | no H2D {no_h2d:.3f} | kernels only {only_k:.3f} | launch sizes {[x[2] for x in sched]} | outs {[(o[0], (o[2]-o[1])//1000) for o in outs]}")
