"""azb_nlist_build_static — the neighbour-list build that allocates nothing (scratch = a static __device__ array inside the
library; include/azb200.h). No counterpart in the reference. Checked through the
C-ABI against the existing entry points (azb_nlist_count + azb_nlist_fill with a caller workspace) on two calls in a row that
share the scratch array: bit-exact arrays (integer work). Collected last on purpose (file name)."""

import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _system(n, seed, bits):
    import azplugins_b200 as az
    import cases

    box, pos, typeid = cases.random_fluid(n, 3.0, 2, seed=seed, jitter=0.5)
    return az.State(az.Snapshot(box, pos, typeid=typeid, types=["A", "B"]), device="cuda", scalar_bits=bits)


def _reference_build(st, r_list):
    """The existing path: azb_nlist_workspace_size / azb_nlist_count / azb_nlist_fill through nlist.Cell."""
    from azplugins_b200 import nlist, pair

    cell = nlist.Cell(buffer=0.4)
    f = pair.DPDConservativeGeneralWeight(nlist=cell, default_r_cut=r_list - 0.4)
    f.r_cut[("B", "B")] = 0.0  # a switched-off type pair
    cell.single_walk = False
    cell.compute(st)
    return cell


def _static_build(lib, st, cell, out, capacity):
    import torch

    from azplugins_b200 import _abi
    from azplugins_b200.state import scalar_dtype

    ftype = scalar_dtype(st.scalar_bits)[0]
    r_list = cell.r_list(st.types)
    rl = r_list.astype(ftype)
    rlsq = torch.from_numpy(np.where(r_list > 0, rl * rl, ftype(-1)).reshape(-1).astype(ftype)).to(st.device)
    a = _abi.NlistArgs()
    a.d_pos, a.N, a.n_max = st.pos.data_ptr(), st.N, st.n_max
    a.box = st.box.as_abi()
    a.r_list_max = float(r_list.max())
    a.d_r_listsq, a.ntypes, a.row_align = rlsq.data_ptr(), len(st.types), 1
    n_neigh, head_list, nl = out
    a.d_n_neigh, a.d_head_list, a.d_nlist, a.capacity = n_neigh.data_ptr(), head_list.data_ptr(), nl.data_ptr(), capacity
    a.d_workspace, a.workspace_bytes = None, 0  # ignored: the scratch is inside the library
    total = C.c_size_t(0)
    stream = C.c_void_p(torch.cuda.current_stream(st.device).cuda_stream)
    rc = lib.azb_nlist_build_static(st.scalar_bits, C.byref(a), C.byref(total), stream)
    torch.cuda.synchronize()
    return rc, int(total.value)


@pytest.mark.parametrize("bits", [32, 64])
def test_static_scratch_list_build_matches_the_workspace_build_twice(bits):
    import torch

    from azplugins_b200 import _abi

    lib = _abi.load()
    # outputs allocated once, up front, by the caller; the entry point itself allocates nothing
    cap = 4_000_000
    out = (torch.zeros(60_000, dtype=torch.int32, device="cuda"), torch.zeros(60_000, dtype=torch.int64, device="cuda"),
           torch.zeros(cap, dtype=torch.int32, device="cuda"))
    # two calls in a row share the scratch array: a larger system first, then a smaller, different one (stale contents of the
    # first call sit behind the second's data)
    for n, seed in ((50_000, 3), (20_000, 4)):
        st = _system(n, seed, bits)
        cell = _reference_build(st, 1.4)
        for t in out:
            t.fill_(-1)
        mem0 = torch.cuda.memory_allocated()
        rc, total = _static_build(lib, st, cell, out, cap)
        assert rc == 0, rc
        assert torch.cuda.memory_allocated() == mem0
        assert total == cell.size_nlist
        assert torch.equal(out[0][:n], cell.n_neigh[:n]) and torch.equal(out[1][:n], cell.head_list[:n])
        assert torch.equal(out[2][:total], cell.nlist[:total])  # same rows in the same (deterministic) order
        assert bool((out[2][total:total + 1000] == -1).all())
    # limits: an output that is too small is reported with the size it takes; a system beyond the scratch array is rejected
    rc, total = _static_build(lib, st, cell, out, 1000)
    assert rc == _abi.ERR_STATIC_SCRATCH_TOO_SMALL and total == cell.size_nlist
    big = _system(600_000, 5, bits)
    cell_big = _reference_build(big, 1.4)
    out_big = (torch.zeros(600_000, dtype=torch.int32, device="cuda"), torch.zeros(600_000, dtype=torch.int64, device="cuda"), out[2])
    rc, _ = _static_build(lib, big, cell_big, out_big, cap)
    assert rc == _abi.ERR_STATIC_SCRATCH_TOO_SMALL
    with pytest.raises(RuntimeError, match="built-in scratch"):
        _abi.check(rc, "azb_nlist_build_static")
