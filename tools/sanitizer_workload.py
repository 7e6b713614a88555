"""A C1-size pass over every kernel family for compute-sanitizer (memcheck / racecheck): neighbour list (two walks, one walk
with rows in place, one walk emitting the tile plans), hash-built plans incl. a pruned one, DPD / PerturbedLJ / ExpandedYukawa / TwoPatchMorse tile and
direct kernels, both bond potentials, NVE steps with a rebuild.
  compute-sanitizer --tool memcheck python tools/sanitizer_workload.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import azplugins_b200 as az
from azplugins_b200 import bond, md, nlist, pair, workloads as W


def main():
    n_launch = 0
    # C5 shape: two forces on one list, pruned view for the DPD force
    snap = W.yukawa_dpd_fluid(n=32000, seed=3)
    st = az.State(snap, device="cuda", scalar_bits=32, seed=42)
    cell = nlist.Cell(buffer=0.4)
    cell.fused_plans = True
    d = pair.DPDGeneralWeight(nlist=cell, kT=1.0, default_r_cut=1.0)
    d.dt = 0.01
    d.params[("A", "A")] = W.C5_DPD_PARAMS[("A", "A")]
    y = pair.ExpandedYukawa(nlist=cell, default_r_cut=2.0)
    y.params[("A", "A")] = W.EY_PARAMS[("A", "A")]
    for f in (d, y):
        f._attach(st)
        f.compute(1)
    cell.invalidate()  # second build: single walk + plans from the list build
    for f in (d, y):
        f.compute(2)
        assert f._plan.from_list_build
        f.use_tiles = False
        f.compute(2)  # direct kernels
        f.use_tiles = True
    ig = md.Integrator(dt=0.01, methods=[md.ConstantVolume()], forces=[d, y]).attach(st)
    ig.run(8)
    n_launch += ig.launches + d.launches + y.launches
    # the default rebuild path: broadcast walk, rows in place from the second build on, hash plan + pruned plan derived from it
    st2 = az.State(W.yukawa_dpd_fluid(n=32000, seed=4), device="cuda", scalar_bits=32, seed=42)
    cell2 = nlist.Cell(buffer=0.4)
    d2 = pair.DPDGeneralWeight(nlist=cell2, kT=1.0, default_r_cut=1.0)
    d2.dt = 0.01
    d2.params[("A", "A")] = W.C5_DPD_PARAMS[("A", "A")]
    y2 = pair.ExpandedYukawa(nlist=cell2, default_r_cut=2.0)
    y2.params[("A", "A")] = W.EY_PARAMS[("A", "A")]
    for f in (d2, y2):
        f._attach(st2)
        f.compute(1)
    cell2.invalidate()
    for f in (d2, y2):
        f.compute(2)
    assert cell2.size_nlist > cell2.n_entries and d2._plan.row_count is not None and not d2._plan.from_list_build
    ig2 = md.Integrator(dt=0.01, methods=[md.ConstantVolume()], forces=[d2, y2]).attach(st2)
    ig2.run(8)
    n_launch += ig2.launches + d2.launches + y2.launches
    # C1 + bonds
    snap = W.polymer_melt(n_chains=300, seed=5)
    st = az.State(snap, device="cuda", scalar_bits=32)
    cell = nlist.Cell(buffer=0.4)
    f = pair.PerturbedLennardJones(nlist=cell, default_r_cut=3.0, mode="shift")
    f.params[("A", "A")] = W.PLJ_PARAMS[("A", "A")]
    q = bond.Quartic()
    q.params["quartic"], q.params["doublewell"] = W.QUARTIC_PARAMS, W.QUARTIC_NEUTRAL
    dw = bond.DoubleWell()
    dw.params["quartic"], dw.params["doublewell"] = W.DOUBLE_WELL_NEUTRAL, W.DOUBLE_WELL_PARAMS
    for g in (f, q, dw):
        g._attach(st)
        g.compute(0)
    q.check_flags()
    dw.check_flags()
    # C4a
    snap = W.patchy_fluid(m=16, seed=7)
    st = az.State(snap, device="cuda", scalar_bits=32)
    cell = nlist.Cell(buffer=0.4)
    t = pair.TwoPatchMorse(nlist=cell, default_r_cut=1.6)
    t.params[("A", "A")] = W.TPM_PARAMS[("A", "A")]
    t._attach(st)
    t.compute(0)
    t.use_tiles = False
    t.compute(0)
    t.use_tiles, t.ring_walk = True, True  # the in-range-compacting walk (opt-in)
    cell.invalidate()
    t.compute(0)
    torch.cuda.synchronize()
    print("sanitizer workload ok", n_launch + f.launches + q.launches + dw.launches + t.launches, "launches")


if __name__ == "__main__":
    main()
