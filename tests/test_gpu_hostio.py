"""The pipelined host step (azplugins_b200/hostio.py) must deliver exactly what the plain launch does."""

import numpy as np
import pytest
import torch

import azplugins_b200 as az
from azplugins_b200 import nlist, pair, workloads as W
from azplugins_b200.hostio import HostStep

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("kind", ["dpd", "plj"])
@pytest.mark.parametrize("chunks", [1, 3, 8])
def test_host_step_matches_plain_launch(kind, chunks):
    if kind == "dpd":
        snap = W.dpd_fluid(n=20000, seed=5)
        nl = nlist.Cell(buffer=0.4)
        f = pair.DPDGeneralWeight(nlist=nl, kT=1.0, default_r_cut=1.0)
        f.dt = 0.01
        for k, v in W.DPD_PARAMS.items():
            f.params[k] = v
    else:
        snap = W.lj_fluid(m=12, seed=6)
        nl = nlist.Cell(buffer=0.4)
        f = pair.PerturbedLennardJones(nlist=nl, default_r_cut=3.0)
        for k, v in W.PLJ_PARAMS.items():
            f.params[k] = v
    st = az.State(snap, device="cuda", scalar_bits=32, seed=42)
    f._attach(st)
    f.compute(3)
    torch.cuda.synchronize()
    want_f = f._force.cpu().clone()
    want_v = f._virial.cpu().clone()
    assert f._plan is not None and f._plan.n_tiles > chunks

    hs = HostStep(f, chunks=chunks)
    h_in, h_out = hs.pinned_inputs(), hs.pinned_outputs()
    for t in st.pos, st.vel:
        t.zero_()  # the step must bring its own inputs
    f._force.zero_()
    hs.run(3, h_in, h_out)
    assert torch.equal(h_out[0], want_f)
    assert torch.equal(f._virial.cpu(), want_v)
    # periodic-boundary tiles read particles from the far end of the index range: they must have been scheduled last
    sched = hs._schedule[-1]
    assert [c[0] for c in sched] == sorted(c[0] for c in sched)
    assert sum(c[2] for c in sched) == f._plan.n_tiles


def test_tile_range_leaves_other_particles_untouched():
    snap = W.dpd_fluid(n=8000, seed=7)
    nl = nlist.Cell(buffer=0.4)
    f = pair.DPDGeneralWeight(nlist=nl, kT=1.0, default_r_cut=1.0)
    f.dt = 0.01
    for k, v in W.DPD_PARAMS.items():
        f.params[k] = v
    st = az.State(snap, device="cuda", scalar_bits=32, seed=42)
    f._attach(st)
    f.compute(0)
    full = f._force.cpu().clone()
    f._force.fill_(-7.0)
    f.compute(0, tile_range=(2, 3))
    got = f._force.cpu()
    ts = f._plan.tile_size
    assert torch.equal(got[2 * ts : 5 * ts], full[2 * ts : 5 * ts])
    assert torch.all(got[: 2 * ts] == -7.0) and torch.all(got[5 * ts :] == -7.0)
    with pytest.raises(RuntimeError):
        f.compute(0, tile_range=(f._plan.n_tiles, 1))


@pytest.mark.parametrize("chunks", [1, 4])
def test_host_step_two_forces_on_one_list(chunks):
    """C5's shape: DPDGeneralWeight + ExpandedYukawa on one neighbour list, evaluated by the launch ranges of one call."""
    snap = W.yukawa_dpd_fluid(n=30000, seed=8)
    nl = nlist.Cell(buffer=0.4)
    d = pair.DPDGeneralWeight(nlist=nl, kT=1.0, default_r_cut=1.0)
    d.dt = 0.01
    for k, v in W.C5_DPD_PARAMS.items():
        d.params[k] = v
    y = pair.ExpandedYukawa(nlist=nl, default_r_cut=2.0)
    for k, v in W.EY_PARAMS.items():
        y.params[k] = v
    st = az.State(snap, device="cuda", scalar_bits=32, seed=42)
    for f in (d, y):
        f._attach(st)
        f.compute(4)
    torch.cuda.synchronize()
    want = [d._force.cpu().clone(), y._force.cpu().clone()]
    hs = HostStep([d, y], chunks=chunks)
    assert len(hs.inputs) == 2 and len(hs.outputs) == 2
    h_in, h_out = hs.pinned_inputs(), hs.pinned_outputs()
    for t in (st.pos, st.vel, d._force, y._force):
        t.zero_()
    hs.run(4, h_in, h_out)
    assert torch.equal(h_out[0], want[0]) and torch.equal(h_out[1], want[1])
    # a list rebuild makes new plans: the cached schedule must not be reused with the old arrays
    nl.invalidate()
    hs.run(5, h_in, h_out)
    d.compute(5)
    y.compute(5)
    torch.cuda.synchronize()
    assert torch.equal(h_out[0], d._force.cpu()) and torch.equal(h_out[1], y._force.cpu())
    hs.release()
