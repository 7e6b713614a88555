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


# ---- migration, ghost re-selection, ghost bonds (SURVEY.md §8 e "on nlist rebuild") -----------------------------------------
def chain_system(n_chains=300, chain_len=20, seed=11):
    """Random-walk chains (bond length 0.5) in a periodic box, random tags; returns box, pos, vel, tag, typeid, bonds (tag pairs)."""
    import azplugins_b200 as az

    rng = np.random.default_rng(seed)
    n = n_chains * chain_len
    L = (n / 3.0) ** (1.0 / 3.0)
    box = az.Box(L * 1.3, L, L / 1.3)
    pos = np.zeros((n_chains, chain_len, 3))
    pos[:, 0] = rng.uniform(-0.5, 0.5, size=(n_chains, 3)) * box.L
    for k in range(1, chain_len):
        step = rng.normal(size=(n_chains, 3))
        pos[:, k] = pos[:, k - 1] + 0.5 * step / np.linalg.norm(step, axis=1, keepdims=True)
    pos = pos.reshape(n, 3)
    pos = (pos + box.L / 2) % box.L - box.L / 2
    tag = rng.permutation(n).astype(np.uint32)
    first = np.arange(n).reshape(n_chains, chain_len)[:, :-1].reshape(-1)
    bonds = np.stack([tag[first], tag[first + 1]], axis=1)
    btype = rng.integers(0, 2, size=len(bonds))
    vel = rng.normal(size=(n, 3))
    typeid = rng.integers(0, 2, size=n)
    return box, pos, vel, tag, typeid, bonds, btype


def displacement_of(tags, it, amp=0.19):
    """Deterministic per-tag displacement, |d| <= amp (a rebuild happens before anything moves half a buffer)."""
    t = tags.astype(np.float64)
    d = np.stack([np.sin(0.37 * t + it), np.cos(0.91 * t + 2 * it), np.sin(1.73 * t + 3 * it + 0.5)], axis=1)
    return amp * d / np.sqrt(3.0)


def run_migration(rank, world, port, device="cpu"):
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", str(port))
    if not dist.is_initialized():
        dist.init_process_group("gloo" if device == "cpu" else "nccl", rank=rank, world_size=world)
    import azplugins_b200 as az
    from azplugins_b200.domain import Decomposition, HaloExchange

    box, pos, vel, tag, typeid, bonds, btype = chain_system()
    n = len(pos)
    dec = Decomposition(world, rank, box)
    own = dec.owns(pos)
    snap = az.Snapshot(box, pos[own], typeid=typeid[own], types=["A", "B"], velocity=vel[own], tag=tag[own])
    dev = device if device == "cpu" else f"cuda:{rank % torch.cuda.device_count()}"
    st = az.State(snap, device=dev, scalar_bits=64)
    if rank == 0:  # the topology may come from any rank(s): it is gathered and de-duplicated
        st.bond_tags, st.bond_typeid = bonds.astype(np.uint32), btype.astype(np.int64)
    halo = HaloExchange(dec, st, 1.4, fields=("pos", "vel"))
    L = box.L
    unwrapped = {int(t): p.copy() for t, p in zip(tag, pos)}  # expected pos + image * L per tag
    vel_by_tag = {int(t): v for t, v in zip(tag, vel)}
    degree = np.bincount(bonds.reshape(-1), minlength=n)

    def check(it):
        bond_len = {}
        for (a, b) in bonds:
            d = unwrapped[int(a)] - unwrapped[int(b)]
            d -= L * np.rint(d / L)
            bond_len[(int(a), int(b))] = bond_len[(int(b), int(a))] = float(np.linalg.norm(d))
        N, ng = st.N, st.n_ghost
        cnt = torch.tensor([N], dtype=torch.int64, device=dev)
        dist.all_reduce(cnt)
        assert int(cnt.item()) == n, (it, int(cnt.item()))
        lp = st.pos[:, :3].cpu().numpy()
        lt = st.tag.cpu().numpy().view(np.uint32)
        assert st.pos.shape[0] == N + ng and len(lt) == N + ng and st.image.shape[0] == N
        assert len(np.unique(lt[:N])) == N
        # locals sit in their own sub-box, carry their velocity, type and image
        assert np.all(dec.owns(lp[:N])), (rank, it)
        img = st.image.cpu().numpy()
        lv = st.vel[:, :3].cpu().numpy()
        ltype = st.pos[:, 3].cpu().numpy().view(np.int64)
        type_by_tag = {int(t): ty for t, ty in zip(tag, typeid)}
        for i in range(N):
            t = int(lt[i])
            assert np.allclose(lp[i] + img[i] * L, unwrapped[t], atol=1e-9), (rank, it, i)
            assert np.allclose(lv[i], vel_by_tag[t]) and int(ltype[i]) == int(type_by_tag[t])
        for g in range(N, N + ng):
            assert np.allclose(lv[g], vel_by_tag[int(lt[g])]) and int(ltype[g]) == int(type_by_tag[int(lt[g])])
        # neighbour sets by tag: local + ghosts in the partly non-periodic local box == global periodic system
        cur = np.array([unwrapped[int(t)] for t in tag])
        cur = (cur + L / 2) % L - L / 2
        got = neighbor_tag_sets(lp, lt, N, L, st.box.periodic, 1.0)
        want = {int(t): s for t, s in zip(tag, neighbor_tag_sets(cur, tag, n, L, (True, True, True), 1.0))}
        for i in range(N):
            assert np.array_equal(got[i], want[int(lt[i])]), (rank, it, i)
        # bonds: every local particle holds all its bonds; the partner (local or ghost) sits at the bond's true length
        tb = st.bond_table()
        nb = tb["n_bonds"].cpu().numpy()
        bl = tb["bondlist"].cpu().numpy().reshape(tb["max_bonds"], tb["pitch"], 2)
        per = np.array(st.box.periodic)
        for i in range(N):
            assert nb[i] == degree[int(lt[i])], (rank, it, i, nb[i], degree[int(lt[i])])
            for s in range(nb[i]):
                j = int(bl[s, i, 0])
                d = lp[i] - lp[j]
                d -= np.where(per, L * np.rint(d / L), 0.0)
                assert abs(np.linalg.norm(d) - bond_len[(int(lt[i]), int(lt[j]))]) < 1e-9, (rank, it, i, j)
        # exclusions follow the same indices
        head, lst = st.exclusions_device()
        assert int(head[-1].item()) == int(nb.sum())

    check(0)
    sent = 0
    for it in range(1, 5):
        lt = st.tag.cpu().numpy().view(np.uint32)
        d_loc = displacement_of(lt[: st.N], it)
        st.pos[: st.N, :3] += torch.from_numpy(d_loc).to(st.pos)
        for t in tag:
            unwrapped[int(t)] = unwrapped[int(t)] + displacement_of(np.array([t]), it)[0]
        halo.rebuild()
        sent = halo.migrated
        check(it)
        # the per-step refresh keeps working on the new selection
        st.vel[: st.N, :3] *= 1.0
        halo.update()
    tot = torch.tensor([sent], dtype=torch.int64, device=dev)
    dist.all_reduce(tot)
    assert int(tot.item()) > 0, "nothing migrated: the test did not exercise the path"
    dist.barrier()
    dist.destroy_process_group()
