"""Time of a neighbour-list rebuild (azb_nlist_count + azb_nlist_fill + tile plan) and of whole NVE steps on the bench
workloads. Usage: python tools/bench_nlist.py [dpd|plj|plj4m]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import azplugins_b200 as az
from azplugins_b200 import md, nlist, pair, workloads


def main():
    which = sys.argv[1] if len(sys.argv) > 1 else "dpd"
    cell = nlist.Cell(buffer=0.4)
    f2 = None
    if which == "eydpd":
        snap = workloads.yukawa_dpd_fluid()
        f = pair.DPDGeneralWeight(nlist=cell, kT=1.0, default_r_cut=1.0)
        f.params[("A", "A")] = workloads.C5_DPD_PARAMS[("A", "A")]
        f2 = pair.ExpandedYukawa(nlist=cell, default_r_cut=2.0)
        f2.params[("A", "A")] = workloads.EY_PARAMS[("A", "A")]
        dt = 0.01
    elif which == "dpd":
        snap = workloads.dpd_fluid()
        f = pair.DPDGeneralWeight(nlist=cell, kT=1.0, default_r_cut=1.0)
        for k, v in workloads.DPD_PARAMS.items():
            f.params[k] = v
        dt = 0.01
    else:
        snap = workloads.lj_fluid(m=100 if which == "plj4m" else 20, jitter=0.05)
        f = pair.PerturbedLennardJones(nlist=cell, default_r_cut=3.0, mode="shift")
        f.params[("A", "A")] = workloads.PLJ_PARAMS[("A", "A")]
        dt = 0.005
    st = az.State(snap, device="cuda", scalar_bits=32, seed=42)
    forces = [f] + ([f2] if f2 is not None else [])
    for g in forces:
        g._attach(st)
        g.compute(0)
    torch.cuda.synchronize()
    for what in ("list", "list+plan"):
        ts = []
        for _ in range(5):
            cell.invalidate()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            cell.compute(st)
            if what == "list+plan":
                for g in forces:
                    g._current_plan()
            torch.cuda.synchronize()
            ts.append(time.perf_counter() - t0)
        print(f"{which}: N={st.N} nbar={cell.mean_neighbors(st.N):.1f} rebuild {what}: min {min(ts) * 1e3:.3f} ms  median {sorted(ts)[2] * 1e3:.3f} ms")
    ig = md.Integrator(dt=dt, methods=[md.ConstantVolume()], forces=forces).attach(st)
    ig.run(20)
    torch.cuda.synchronize()
    r0 = ig.nlist_rebuilds
    t0 = time.perf_counter()
    steps = 200
    ig.run(steps)
    torch.cuda.synchronize()
    t = time.perf_counter() - t0
    print(f"{which}: NVE {steps} steps in {t * 1e3:.1f} ms = {steps / t:.1f} steps/s = {st.N * steps / t / 1e9:.3f} G particle-steps/s, "
          f"{ig.nlist_rebuilds - r0} list rebuilds, E={ig.kinetic_energy + ig.potential_energy:.3f}")


if __name__ == "__main__":
    main()
