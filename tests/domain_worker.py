"""Worker for the multi-process domain-decomposition tests (gloo on CPU; also run under torchrun on GPUs)."""

import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def global_system(n=6000, rho=3.0, seed=5):
    import azplugins_b200 as az

    rng = np.random.default_rng(seed)
    L = (n / rho) ** (1.0 / 3.0)
    box = az.Box(L * 1.3, L, L / 1.3)
    pos = rng.uniform(-0.5, 0.5, size=(n, 3)) * box.L
    vel = rng.normal(size=(n, 3))
    tag = rng.permutation(n).astype(np.uint32)
    typeid = rng.integers(0, 2, size=n)
    return box, pos, vel, tag, typeid


def neighbor_tag_sets(pos, tags, n_local, box_L, periodic, r):
    """Brute-force: for each of the first n_local particles the sorted tags of all particles within r."""
    out = []
    for i in range(n_local):
        d = pos - pos[i]
        for k in range(3):
            if periodic[k]:
                d[:, k] -= box_L[k] * np.rint(d[:, k] / box_L[k])
        m = (np.sum(d * d, axis=1) < r * r)
        m[i] = False
        out.append(np.sort(tags[m]))
    return out


def run(rank, world, port, device="cpu", r_ghost=1.4, result=None):
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", str(port))
    if not dist.is_initialized():
        dist.init_process_group("gloo" if device == "cpu" else "nccl", rank=rank, world_size=world)
    import azplugins_b200 as az
    from azplugins_b200.domain import Decomposition, HaloExchange

    box, pos, vel, tag, typeid = global_system()
    dec = Decomposition(world, rank, box)
    own = dec.owns(pos)
    snap = az.Snapshot(box, pos[own], typeid=typeid[own], types=["A", "B"], velocity=vel[own], tag=tag[own])
    dev = device if device == "cpu" else f"cuda:{rank % torch.cuda.device_count()}"
    st = az.State(snap, device=dev, scalar_bits=64)
    halo = HaloExchange(dec, st, r_ghost, fields=("pos", "vel"))
    # every particle is owned exactly once
    cnt = torch.tensor([st.N], dtype=torch.int64, device=dev)
    dist.all_reduce(cnt)
    assert int(cnt.item()) == len(pos)
    assert st.n_ghost > 0 and st.pos.shape[0] == st.N + st.n_ghost
    lb = st.box
    assert lb.periodic == tuple(g == 1 for g in dec.grid)

    # (1) neighbour sets by tag: local + ghosts in the local (partly non-periodic) box == global periodic system
    lp = st.pos[:, :3].cpu().numpy()
    lt = st.tag.cpu().numpy().view(np.uint32)
    got = neighbor_tag_sets(lp, lt, st.N, box.L, lb.periodic, 1.0)
    want_all = neighbor_tag_sets(pos, tag, len(pos), box.L, (True, True, True), 1.0)
    by_tag = {int(t): s for t, s in zip(tag, want_all)}
    for i in range(st.N):
        assert np.array_equal(got[i], by_tag[int(lt[i])]), (rank, i)
    # (2) ghosts carry their owner's type and velocity
    vel_by_tag = {int(t): v for t, v in zip(tag, vel)}
    type_by_tag = {int(t): ty for t, ty in zip(tag, typeid)}
    gv = st.vel[st.N:, :3].cpu().numpy()
    gtype = st.pos[st.N:, 3].cpu().numpy().view(np.int64)
    for g in range(st.n_ghost):
        assert np.allclose(gv[g], vel_by_tag[int(lt[st.N + g])])
        assert int(gtype[g]) == int(type_by_tag[int(lt[st.N + g])])
    # (3) per-step update: move the locals, refresh, ghosts follow their owners (shift included)
    disp = 0.01 * np.sin(np.arange(len(pos)))[:, None] * np.ones(3)
    disp_by_tag = {int(t): d for t, d in zip(tag, disp)}
    before = st.pos[st.N:, :3].clone()
    d_loc = np.stack([disp_by_tag[int(t)] for t in lt[: st.N]])
    st.pos[: st.N, :3] += torch.from_numpy(d_loc).to(st.pos)
    halo.update()
    d_gh = np.stack([disp_by_tag[int(t)] for t in lt[st.N:]])
    assert torch.allclose(st.pos[st.N:, :3], before + torch.from_numpy(d_gh).to(before), atol=1e-12)
    assert halo.bytes_per_update() == sum(int(i.numel()) for i in halo.send_idx) * 32 * 2
    dist.barrier()
    if result is not None:
        result[rank] = (st.N, st.n_ghost)
    dist.destroy_process_group()


if __name__ == "__main__":
    run(int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("MASTER_PORT", "29511")),
        device="cuda" if torch.cuda.is_available() else "cpu")
    print("domain worker ok", os.environ["RANK"])
