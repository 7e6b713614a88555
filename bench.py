"""bench.py — force-compute throughput of the B200 engine (and, with --impl reference, of the reference's CPU path).

One "step" = one evaluation of the hot path (the force kernel launches through the C-ABI) over the whole synthetic
configuration. Default workload: BASELINE.json configs[4] at its single-GPU size — C5, ExpandedYukawa r_cut = 2.0 +
DPDGeneralWeight r_cut = 1.0 on one neighbour list, 8 M particles per GPU (weak scaling: 8 M -> 64 M on 8 GPUs),
Scalar = float. It is the largest single-GPU configuration of BASELINE.json and the one its scaling target is quoted on.
Rank 0 prints ONE JSON line; at N = 1 the line also carries the same measurement in double precision ("f64"), the other
configurations ("others": C2 / 4 M LJ / C3 melt / C4a), the cost of the list + tile-plan rebuild amortised over a measured
rebuild period ("amortised") and the reference CPU path on the host cores ("cpu_baseline").

  python bench.py --gpus 1 --steps 20 --warmup 5
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...
  python bench.py --impl reference            (CPU: reference evaluator headers + restated HOOMD loop, all host threads)
"""

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ALL_CPUS = os.sched_getaffinity(0)  # the CPU baseline always runs on every core the process was given
METRIC = "force-compute particle-steps/s"
UNIT = "particle-steps/s"
WORKLOADS = ["eydpd", "dpd", "plj", "plj4m", "patchy", "hertz", "colloid", "melt"]
DEFAULT_N = {"dpd": 1_000_188, "plj": 32_000, "plj4m": 4_000_000, "patchy": 2_048_000, "hertz": 2_048_000, "colloid": 1_000_000,
             "melt": 4_000_000, "eydpd": 8_000_000}
L2_NOTE = "GPU arm: flushed between timed steps (384 MiB fill); reference arm: host caches, not flushed"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="eydpd", choices=WORKLOADS)
    ap.add_argument("--n", type=int, default=0, help="override particle count (per GPU)")
    ap.add_argument("--strong", action="store_true", help="strong scaling: the workload's particle count is the TOTAL, split over the GPUs")
    ap.add_argument("--bits", type=int, default=32, choices=[32, 64])
    ap.add_argument("--tpp", type=int, default=0)
    ap.add_argument("--block", type=int, default=0)
    ap.add_argument("--no-tiles", action="store_true", help="gather neighbour data from global memory instead of per-tile shared-memory staging")
    ap.add_argument("--tile-size", type=int, default=0, help="particles per tile of the tile plan (0 = the force's own choice)")
    ap.add_argument("--no-virial", action="store_true")
    ap.add_argument("--no-flush", action="store_true", help="do not flush L2 between timed steps")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the f64 / others / amortised / strong sub-records")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--min-seconds", type=float, default=0.6, help="the K-step block is repeated until the timed region is this long (clock sampling)")
    ap.add_argument("--e2e-chunks", type=int, default=8, help="chunks of the pipelined host step (1 = copy, compute, copy)")
    ap.add_argument("--e2e-taper", default="8,4,2", help="shares (percent of the rows) of the LAST input chunks of the pipelined host step, "
                    "e.g. 8,4,2; 'none' = equal chunks (hostio.chunk_edges)")
    return ap.parse_args()


def bind_to_gpu_numa_node(index):
    """Run this process (and so first-touch its pinned host buffers) on the CPUs local to GPU `index`: host<->device
    copies from the far NUMA node lose a third or more of the PCIe bandwidth. Best effort; returns a note for the JSON line."""
    try:
        import pynvml

        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        bus = pynvml.nvmlDeviceGetPciInfo(h).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        dom, rest = bus.split(":", 1)
        path = f"/sys/bus/pci/devices/{dom[-4:].lower()}:{rest.lower()}"
        cpus = open(path + "/local_cpulist").read().strip()
        node = open(path + "/numa_node").read().strip()
        ids = set()
        for part in cpus.split(","):
            lo, _, hi = part.partition("-")
            ids.update(range(int(lo), int(hi or lo) + 1))
        ids &= os.sched_getaffinity(0)
        if ids:
            os.sched_setaffinity(0, ids)
            if node in ("-1", ""):
                return f"host exposes a single NUMA domain (node {node or '?'}, cpus {cpus}): pinned buffers cannot be placed closer to GPU {bus}"
            return f"process bound to NUMA node {node} (cpus {cpus}) of GPU {bus}"
        return f"GPU {bus}: no usable local cpus ({cpus})"
    except Exception as exc:  # noqa: BLE001 - affinity is an optimisation, never a failure
        return f"not bound ({type(exc).__name__})"


# ---------------------------------------------------------------------------------------------------------------
def make_workload(name, n_override, rank=0):
    """Returns (label, snapshot, setup, buffer) where setup(nlist) -> list of force objects (first = dominant kernel)."""
    from azplugins_b200 import bond, pair, workloads as W

    if name == "dpd":
        n = n_override or DEFAULT_N["dpd"]
        snap = W.dpd_fluid(n=n, seed=2001 + 17 * rank)

        def setup(nl):
            f = pair.DPDGeneralWeight(nlist=nl, kT=1.0, default_r_cut=1.0)
            f.dt = 0.01
            for k, v in W.DPD_PARAMS.items():
                f.params[k] = v
            return [f]

        return f"C2 DPDGeneralWeight fluid N={n} rho=3 r_cut=1.0 buffer=0.4 s=2 kT=1 dt=0.01 seed=42", snap, setup, 0.4
    if name in ("plj", "plj4m"):
        m = 20 if name == "plj" else 100
        if n_override:
            m = max(2, round((n_override / 4) ** (1 / 3)))
        snap = W.lj_fluid(m=m, seed=1001 + 17 * rank)

        def setup(nl):
            f = pair.PerturbedLennardJones(nlist=nl, default_r_cut=3.0)
            for k, v in W.PLJ_PARAMS.items():
                f.params[k] = v
            return [f]

        return f"C1-style PerturbedLennardJones fluid N={snap.N} rho=0.8 r_cut=3.0 buffer=0.4 lambda=0.5", snap, setup, 0.4
    if name == "patchy":
        m = 80 if not n_override else max(2, round((n_override / 4) ** (1 / 3)))
        snap = W.patchy_fluid(m=m, seed=4001 + 17 * rank)

        def setup(nl):
            f = pair.TwoPatchMorse(nlist=nl, default_r_cut=1.6)
            for k, v in W.TPM_PARAMS.items():
                f.params[k] = v
            return [f]

        return f"C4a TwoPatchMorse N={snap.N} rho=0.6 r_cut=1.6 buffer=0.4", snap, setup, 0.4
    if name == "hertz":
        m = 80 if not n_override else max(2, round((n_override / 4) ** (1 / 3)))
        snap = W.hertz_mixture(m=m, seed=4101 + 17 * rank)

        def setup(nl):
            f = pair.Hertz(nlist=nl, default_r_cut=0.0)
            par, rc = W.hertz_tables()
            for k, v in par.items():
                f.params[k] = v
                f.r_cut[k] = rc[k]
            return [f]

        return f"C4b polydisperse Hertz mixture N={snap.N} rho=1.2, 4 species, r_cut = mean diameter (0.85..1.15), buffer=0.4", snap, setup, 0.4
    if name == "colloid":
        m = 75 if not n_override else max(2, round((n_override / 4 / 0.59) ** (1 / 3)))
        snap = W.colloid_mixture(m=m, seed=4201 + 17 * rank)

        def setup(nl):
            f = pair.Colloid(nlist=nl, default_r_cut=0.0)
            for k, v in W.COLLOID_PARAMS.items():
                f.params[k] = v
                f.r_cut[k] = W.COLLOID_RCUT[k]
            return [f]

        return (f"C4b Colloid mixture N={snap.N}: colloids of radius 1.5 on a cubic lattice in a solvent at rho=0.6, "
                "r_cut 3.0 / 4.5 / 6.0 (solvent-solvent / colloid-solvent / colloid-colloid), buffer=0.4"), snap, setup, 0.4
    if name == "melt":
        chains = 40_000 if not n_override else max(1, n_override // 100)
        snap = W.polymer_melt(n_chains=chains, seed=3001 + 17 * rank)

        def setup(nl):
            f = pair.PerturbedLennardJones(nlist=nl, default_r_cut=3.0)
            for k, v in W.PLJ_PARAMS.items():
                f.params[k] = v
            # one bond object per potential, as in HOOMD; each is neutral (zero force and energy) on the other's bond type
            q = bond.Quartic()
            q.params["quartic"] = W.QUARTIC_PARAMS
            q.params["doublewell"] = W.QUARTIC_NEUTRAL
            dw = bond.DoubleWell()
            dw.params["quartic"] = W.DOUBLE_WELL_NEUTRAL
            dw.params["doublewell"] = W.DOUBLE_WELL_PARAMS
            return [f, q, dw]

        return (f"C3 bead-spring melt N={snap.N} (chains of 100) PerturbedLJ r_cut=3.0 (bonded pairs excluded) + bond.Quartic (90% of bonds) "
                "+ bond.DoubleWell (10%)"), snap, setup, 0.4
    if name == "eydpd":
        n = n_override or DEFAULT_N["eydpd"]
        snap = W.yukawa_dpd_fluid(n=n, seed=5001 + 17 * rank)

        def setup(nl):
            d = pair.DPDGeneralWeight(nlist=nl, kT=1.0, default_r_cut=1.0)
            d.dt = 0.01
            for k, v in W.C5_DPD_PARAMS.items():
                d.params[k] = v
            y = pair.ExpandedYukawa(nlist=nl, default_r_cut=2.0)
            for k, v in W.EY_PARAMS.items():
                y.params[k] = v
            return [d, y]

        return (f"C5 DPDGeneralWeight r_cut=1.0 + ExpandedYukawa r_cut=2.0 on one neighbour list, N={snap.N} rho=1.0 buffer=0.4 "
                "(roofline: the slower of the two kernels)"), snap, setup, 0.4
    raise ValueError(name)


def list_cutoff(workload, snap):
    """r_cut (scalar or per-type-pair matrix) the workload's neighbour list is built for."""
    from azplugins_b200 import workloads as W

    if workload in ("hertz", "colloid"):
        by_pair = W.hertz_tables()[1] if workload == "hertz" else W.COLLOID_RCUT
        names = snap.types
        return np.array([[by_pair[tuple(sorted((a, b)))] for b in names] for a in names])
    return {"dpd": 1.0, "plj": 3.0, "plj4m": 3.0, "melt": 3.0, "patchy": 1.6, "eydpd": 2.0}[workload]


def algorithmic_bytes_per_particle(force, nbar, bits, virial):
    """SURVEY.md §8 d: compulsory HBM bytes per particle-step of one kernel (nbar = mean list length; for a bond force the
    mean number of bond-table slots per particle)."""
    from azplugins_b200 import _abi

    s = bits // 8
    S4 = 4 * s
    v = 6 * s if virial else 0
    if force._kind == _abi.KIND_BOND:
        return S4 + 4 + S4 + v + 12 * nbar
    b = S4 + 4 + 8 + S4 + v + 4 * nbar
    if force._kind == _abi.KIND_DPD:
        b += S4 + 4
    if force._kind == _abi.KIND_ANISO:
        b += 2 * S4
    return b


def shared_config(label, n, virial, world, bits, strong=False):
    """`config` of the JSON line: the description of the WORKLOAD, identical in both arms (the driver compares them);
    what is specific to an arm (tile plan, halo transport, CPU threads) sits in other keys of the line."""
    from azplugins_b200.domain import GRIDS

    par = "1 GPU" if world == 1 else (f"{world} GPUs, spatial domains {GRIDS[world]}, ghost halo refreshed every step, "
                                      + ("strong scaling (total fixed)" if strong else "weak scaling (per-GPU size fixed)"))
    return {"workload": label, "particles_per_gpu": int(n), "compute_virial": bool(virial), "scalar": "float" if bits == 32 else "double",
            "neighbor_list": "full, HOOMD layout (n_neigh / head_list / nlist), r_cut + buffer", "l2": L2_NOTE, "parallelism": par}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
            time.sleep(0.25)  # nvidia-smi needs a moment before its first sample
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.perf_counter(), line.strip()))

    def mark(self):
        return time.perf_counter()

    def stop(self, t_begin=None, t_end=None):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "samples": 0, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.12)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for when, ln in self.lines:
            if t_begin is not None and not (t_begin <= when <= t_end + 0.06):
                continue
            p = [x.strip() for x in ln.split(",")]
            if len(p) < 7:
                continue
            try:
                sm.append(float(p[0]))
                mx.append(float(p[1]))
            except ValueError:
                continue
            for nm, val in zip(names, p[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": float(np.max(mx)) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------------------------------------
def host_arrays_for_oracle(snap, bits, n_neigh, head_list, nlist):
    """HostSystem (tests/oracle_api.py) filled from a Snapshot and HOOMD-layout neighbour arrays."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_api as O

    hs = O.HostSystem(snap.box, snap.position, snap.typeid, len(snap.types), bits, velocity=snap.velocity, mass=snap.mass,
                      orientation=snap.orientation, tag=snap.tag)
    hs.n_neigh, hs.head_list, hs.nlist = n_neigh, head_list, nlist
    return hs, O


def build_checkers():
    """Compile the CPU oracles if they are missing (make; never touches or loads the CUDA library)."""
    port = os.path.join(ROOT, "oracle", "_build", "libazoracle.so")
    ref = os.path.join(ROOT, "oracle", "_ref", "libazref64.so")
    if not os.path.exists(port):
        subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "port"], check=True)
    if os.path.isdir("/root/reference/src") and not os.path.exists(ref):
        subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "ref"], check=True)


def reference_step_fn(workload, snap, bits, nl_arrays, nthreads, half_list=False):
    """step(t) evaluating the workload's forces with the reference's own CPU implementation of the path: its evaluator headers
    compiled unmodified (oracle/_ref) — or the C port when /root/reference was not available at build time — inside the
    restated HOOMD host loops. Full list + OpenMP over particles, or (half_list) HOOMD's CPU mode: half list + third law."""
    build_checkers()
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import cases
    import oracle_api as O
    from azplugins_b200 import workloads as W
    from azplugins_b200.state import build_bond_table

    hs, _ = host_arrays_for_oracle(snap, bits, *nl_arrays)
    if half_list:
        hs = half_list_of(hs)
    which = "ref" if O.have_ref() else "port"
    T = len(snap.types)
    names = snap.types
    kw = dict(nthreads=nthreads, which=which, half_list=half_list)

    def tables(spec, pdict, rc):
        par_by = {(min(i, j), max(i, j)): pdict[tuple(sorted((names[i], names[j])))] for i in range(T) for j in range(T)}
        rc_by = {(min(i, j), max(i, j)): (rc[tuple(sorted((names[i], names[j])))] if isinstance(rc, dict) else rc) for i in range(T) for j in range(T)}
        return cases.pair_tables(spec, par_by, rc_by, T, bits)

    if workload == "dpd":
        par, rcsq, _, _ = tables(cases.SPEC_BY_API["DPDGeneralWeight"], W.DPD_PARAMS, 1.0)
        step = lambda t: hs.dpd(par, rcsq, seed=42, timestep=t, dt=0.01, kT=1.0, **kw)  # noqa: E731
    elif workload in ("plj", "plj4m"):
        spec = cases.SPEC_BY_API["PerturbedLennardJones"]
        par, rcsq, ronsq, _ = tables(spec, W.PLJ_PARAMS, 3.0)
        step = lambda t: hs.pair(spec.potential_id, par, rcsq, ronsq, **kw)  # noqa: E731
    elif workload == "melt":
        spec = cases.SPEC_BY_API["PerturbedLennardJones"]
        par, rcsq, ronsq, _ = tables(spec, W.PLJ_PARAMS, 3.0)
        table = build_bond_table(snap.N, snap.N, snap.bond_group, snap.bond_typeid)
        q, dw = cases.SPEC_BY_API["Quartic"], cases.SPEC_BY_API["DoubleWell"]
        par_q = np.array([q.pack(W.QUARTIC_PARAMS, bits), q.pack(W.QUARTIC_NEUTRAL, bits)])
        par_dw = np.array([dw.pack(W.DOUBLE_WELL_NEUTRAL, bits), dw.pack(W.DOUBLE_WELL_PARAMS, bits)])

        def step(t):  # pair + both bond objects: what the GPU arm evaluates
            hs.pair(spec.potential_id, par, rcsq, ronsq, **kw)
            hs.bond(q.potential_id, par_q, *table, 2, nthreads=nthreads, which=which)
            hs.bond(dw.potential_id, par_dw, *table, 2, nthreads=nthreads, which=which)
    elif workload in ("hertz", "colloid"):
        spec = cases.SPEC_BY_API["Hertz" if workload == "hertz" else "Colloid"]
        pdict, rc = W.hertz_tables() if workload == "hertz" else (W.COLLOID_PARAMS, W.COLLOID_RCUT)
        par, rcsq, ronsq, _ = tables(spec, pdict, rc)
        step = lambda t: hs.pair(spec.potential_id, par, rcsq, ronsq, **kw)  # noqa: E731
    elif workload == "patchy":
        par, rcsq, _, _ = tables(cases.SPEC_BY_API["TwoPatchMorse"], W.TPM_PARAMS, 1.6)
        step = lambda t: hs.aniso(par, rcsq, **kw)  # noqa: E731
    elif workload == "eydpd":
        par_d, rcsq_d, _, _ = tables(cases.SPEC_BY_API["DPDGeneralWeight"], W.C5_DPD_PARAMS, 1.0)
        spec_y = cases.SPEC_BY_API["ExpandedYukawa"]
        par_y, rcsq_y, ronsq_y, _ = tables(spec_y, W.EY_PARAMS, 2.0)

        def step(t):
            hs.dpd(par_d, rcsq_d, seed=42, timestep=t, dt=0.01, kT=1.0, **kw)
            hs.pair(spec_y.potential_id, par_y, rcsq_y, ronsq_y, **kw)
    else:
        raise ValueError(workload)
    return step, hs.N, ("reference" if which == "ref" else "port")


def half_list_of(hs):
    """Half-storage copy of a full list (pairs i < j once), vectorised: HOOMD's CPU neighbour-list mode."""
    N = hs.N
    nn = hs.n_neigh[:N].astype(np.int64)
    starts = hs.head_list[:N].astype(np.int64)
    i_idx = np.repeat(np.arange(N), nn)
    offs = np.arange(len(i_idx)) - np.repeat(np.cumsum(nn) - nn, nn)
    j_idx = hs.nlist[np.repeat(starts, nn) + offs]
    keep = (j_idx > i_idx) | (j_idx >= N)
    half = type(hs).__new__(type(hs))
    half.__dict__.update(hs.__dict__)
    cnt = np.bincount(i_idx[keep], minlength=N).astype(np.uint32)
    head = np.zeros(N, dtype=np.uint64)
    head[1:] = np.cumsum(cnt[:-1])
    half.n_neigh, half.head_list, half.nlist = cnt, head, np.ascontiguousarray(j_idx[keep], dtype=np.uint32)
    half.force, half.torque, half.virial = np.zeros_like(hs.force), np.zeros_like(hs.torque), np.zeros_like(hs.virial)
    return half


def time_reference_cpu(workload, snap, bits, nl_arrays, seconds, steps=None, warmup=1, nthreads=None, half_list=False):
    """Returns (particle-steps/s, cores, kind, seconds/step, repetitions)."""
    cores = nthreads or len(ALL_CPUS)
    os.sched_setaffinity(0, ALL_CPUS)  # undo the GPU-local binding of the copy path: the reference gets all host cores
    step, n, kind = reference_step_fn(workload, snap, bits, nl_arrays, cores, half_list)
    for t in range(warmup):
        step(t)
    times = []
    t_begin = time.perf_counter()
    k = 0
    while True:
        t0 = time.perf_counter()
        step(k)
        times.append(time.perf_counter() - t0)
        k += 1
        if steps is not None:
            if k >= steps:
                break
        elif k >= 2 and time.perf_counter() - t_begin > seconds:
            break
    mean = float(np.mean(times))
    return n / mean, cores, kind, mean, len(times)


def oracle_nlist(snap, bits, r_list, exclusions=None):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_api as O

    hs = O.HostSystem(snap.box, snap.position, snap.typeid, len(snap.types), bits)
    if exclusions is not None:
        hs.build_nlist(r_list, excl_head=exclusions[0], excl_list=exclusions[1])
    else:
        hs.build_nlist(r_list)
    return hs.n_neigh, hs.head_list, hs.nlist


def bond_exclusions(snap):
    """CSR of the bonded partners of every particle (HOOMD nlist default exclusions=('bond',)), host arrays."""
    g = snap.bond_group
    if not len(g):
        return None
    a = np.concatenate([g[:, 0], g[:, 1]])
    b = np.concatenate([g[:, 1], g[:, 0]])
    order = np.argsort(a, kind="stable")
    a, b = a[order], b[order]
    head = np.zeros(snap.N + 1, dtype=np.uint64)
    np.cumsum(np.bincount(a, minlength=snap.N)[: snap.N], out=head[1:])
    return head, b.astype(np.uint32)


def run_reference(args):
    """--impl reference: the reference's CPU path on this box's host cores (rank 0 only). No CUDA library is loaded."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = int(os.environ.get("WORLD_SIZE", "1"))
    label, snap, _, buffer = make_workload(args.workload, args.n)
    rc = list_cutoff(args.workload, snap)
    t0 = time.perf_counter()
    nl = oracle_nlist(snap, args.bits, rc + buffer, bond_exclusions(snap))  # input preparation, not timed
    prep = time.perf_counter() - t0
    value, cores, kind, sec, k = time_reference_cpu(args.workload, snap, args.bits, nl, 0.0, steps=args.steps, warmup=args.warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32" if args.bits == 32 else "f64", "data": "synthetic",
        "config": shared_config(label, snap.N, not args.no_virial, world, args.bits),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind,
                         "sample": f"every step = the whole workload ({snap.N} particles) on rank 0's host, OpenMP over particles, full neighbour list "
                                   f"(built by the CPU cell-list oracle in {prep:.1f} s, not timed)"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


# ---------------------------------------------------------------------------------------------------------------
class Rig:
    """One workload resident on this rank's GPU: state, neighbour list, force objects (+ halo exchange at N > 1)."""

    def __init__(self, args, workload, bits, dev, world, rank, n_per_gpu=0, strong=False):
        import azplugins_b200 as az
        from azplugins_b200 import _abi, nlist

        self.args, self.workload, self.bits, self.dev, self.world, self.rank = args, workload, bits, dev, world, rank
        self.label, snap, setup, buffer = make_workload(workload, n_per_gpu, rank)
        self.snap = snap
        self.nl = nlist.Cell(buffer=buffer)
        self.forces = setup(self.nl)
        self.halo = None
        self.strong = strong
        dec = None
        if world > 1:
            # every rank generates its own sub-domain (same recipe, rank-dependent seed) of one periodic global box of
            # grid x sub-box; ghosts within r_cut + buffer are exchanged with the neighbouring domains every step
            from azplugins_b200.domain import GRIDS, Decomposition

            grid = GRIDS[world]
            sub = snap.box
            gbox = az.Box(sub.Lx * grid[0], sub.Ly * grid[1], sub.Lz * grid[2])
            dec = Decomposition(world, rank, gbox)
            snap.position += (dec.lo + dec.hi) / 2
            snap.tag = (snap.tag.astype(np.uint64) + rank * snap.N).astype(np.uint32)
            snap.box = gbox
        self.st = st = az.State(snap, device=dev, scalar_bits=bits, seed=42)
        if world > 1:
            from azplugins_b200.domain import HaloExchange

            kinds = {f._kind for f in self.forces}
            fields = ("pos",) + (("vel",) if _abi.KIND_DPD in kinds else ()) + (("orientation",) if _abi.KIND_ANISO in kinds else ())
            self.halo = HaloExchange(dec, st, float(self.nl.r_list(st.types).max()), fields=fields)
        self.virial = not args.no_virial
        for f in self.forces:
            f.compute_virial = self.virial
            f.threads_per_particle = args.tpp
            f.block_size = args.block
            f.use_tiles = not args.no_tiles
            f.tile_size = args.tile_size
            f._attach(st)
        import torch

        t0 = time.perf_counter()
        self.nl.compute(st)
        torch.cuda.synchronize()
        self.nlist_build_s = time.perf_counter() - t0
        self.nbar = self.nl.mean_neighbors(st.N)

    def launches(self):
        return sum(f.launches for f in self.forces) + (self.halo.launches if self.halo is not None else 0)

    def step(self, t):
        if self.halo is not None:
            self.halo.step(self.forces[0], t)  # ghost refresh overlapped with the ghost-free tiles (azplugins_b200/domain.py)
            for f in self.forces[1:]:
                f.compute(t)
            return
        for f in self.forces:
            f.compute(t)

    def barrier(self):
        import torch
        import torch.distributed as dist

        if self.world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def time_steps(self, K, t_first, flush, min_seconds=0.0, max_blocks=64):
        """Times blocks of exactly K steps (CUDA events on the launching stream around every step and around every force
        launch; L2 flushed between steps). The block is repeated until `min_seconds` have passed so that the clock sampler
        sees the region; returns per-block step times [R, K] and per-force launch times [R, K, F] in ms."""
        import torch

        step_ms, kern_ms = [], []
        t_begin = time.perf_counter()
        for r in range(max_blocks):
            ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
            kev = [[(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in self.forces] for _ in range(K)]
            self.barrier()
            for k in range(K):
                if flush is not None:
                    flush.fill_(k & 0xFF)
                ev[k][0].record()
                for fi, f in enumerate(self.forces):
                    kev[k][fi][0].record()
                    if self.halo is not None and fi == 0:
                        self.halo.step(f, t_first + k)
                    else:
                        f.compute(t_first + k)
                    kev[k][fi][1].record()
                ev[k][1].record()
            self.barrier()
            step_ms.append([a.elapsed_time(b) for a, b in ev])
            kern_ms.append([[a.elapsed_time(b) for a, b in row] for row in kev])
            done = time.perf_counter() - t_begin >= min_seconds
            if self.world > 1:
                import torch.distributed as dist

                flag = torch.tensor([1 if done else 0], device=self.dev)
                dist.all_reduce(flag, op=dist.ReduceOp.MIN)  # every rank runs the same number of blocks
                done = bool(flag.item())
            if done:
                break
        return np.array(step_ms), np.array(kern_ms)

    def kernel_records(self, kern_ms, peak):
        """Per force object: mean launch time, algorithmic bytes per particle (SURVEY.md §8 d), achieved GB/s and fraction."""
        from azplugins_b200 import _abi

        out = []
        st = self.st
        for fi, f in enumerate(self.forces):
            if f._kind == _abi.KIND_BOND:
                nbar = float(st.bond_table()["n_bonds"].sum().item()) / max(st.N, 1)
            else:
                nbar = self.nbar
            plan = getattr(f, "_plan", None)
            pruned = plan is not None and plan.row_count is not None
            if pruned:
                # a force walking a pruned view of the list (azb_tile_build_args.d_prune_pos) is charged for the entries its view
                # holds, not for the list's: the conservative reading of SURVEY.md §8 d's "4 n" term
                nbar = plan.entries / max(st.N, 1)
            bpp = algorithmic_bytes_per_particle(f, nbar, self.bits, self.virial)
            ms = float(kern_ms[..., fi].mean())
            ach = st.N * bpp / (ms * 1e-3) / 1e9
            rec = {"kernel": type(f).__name__, "kernel_ms": ms, "bytes_per_particle": bpp, "achieved": ach, "frac": ach / peak,
                   "entries_per_particle": nbar}
            if pruned:
                full = algorithmic_bytes_per_particle(f, self.nbar, self.bits, self.virial)
                rec["pruned_view"] = {"list_entries_per_particle": self.nbar, "bytes_per_particle_full_list": full,
                                      "frac_full_list": st.N * full / (ms * 1e-3) / 1e9 / peak}
            out.append(rec)
        return out


def hbm_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        return float(json.load(open(path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (measured copy, burst)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def roofs_from_profile(workload, bits, roofline, kernel):
    """The other two roofs SURVEY.md §8 d asks for — instruction issue and the L1 / shared-memory data pipe — cannot be measured
    from inside the process; they are taken from the committed ncu capture of this workload's dominant kernel
    (profiles/roofs_<workload>_f<bits>.json, written by tools/ncu_roofs.py from an `ncu --set full` report). `bound` names the
    largest of the three utilisations."""
    path = os.path.join(ROOT, "profiles", f"roofs_{workload}_f{bits}.json")
    fracs = {"hbm": roofline["frac"]}
    r = json.load(open(path)) if os.path.exists(path) else None
    if r is not None and r.get("kernel_class") not in (None, kernel):
        r = None  # the capture is of another force object of this workload
    if r is not None:
        roofline["traffic"] = r.get("dram_bytes_per_launch")
        roofline["issue_frac"] = r.get("issue_active_pct", 0.0) / 100.0
        roofline["l1_smem_frac"] = r.get("l1_data_pipe_pct", 0.0) / 100.0
        roofline["dram_frac_ncu"] = r.get("dram_pct", 0.0) / 100.0
        roofline["roofs_source"] = r.get("source")
        fracs["issue"] = roofline["issue_frac"]
        fracs["l1_smem"] = roofline["l1_smem_frac"]
        # cycles the FMA pipe is held (an FFMA2 holds it for two and counts once as an instruction) and the XU pipe (MUFU: 16 lanes
        # per clock and SM, tools/microbench/mufu.cu), when the capture has them
        if r.get("fma_pipe_cycles_pct") is not None:
            roofline["fma_pipe_frac"] = fracs["fma_pipe"] = r["fma_pipe_cycles_pct"] / 100.0
        if r.get("xu_pipe_pct") is not None:
            roofline["xu_pipe_frac"] = fracs["xu_pipe"] = r["xu_pipe_pct"] / 100.0
    else:
        roofline["issue_frac"] = roofline["l1_smem_frac"] = None
        roofline["roofs_source"] = "no ncu capture committed for this workload / precision"
    # the contract's enum is hbm | tensor; tensor cores are not on this path. `bound` stays "hbm" (the roof `frac` is measured
    # against) and `binding_roof` names the unit that actually limits the kernel.
    roofline["bound"] = "hbm"
    roofline["binding_roof"] = max(fracs, key=fracs.get)
    return roofline


def mgpu_check(rig):
    """Correctness of the decomposed evaluation, checked where the driver's scaling run sees it: (1) Newton's third law across
    ALL ranks (the full list evaluates every pair on both owners; a wrong or stale ghost row breaks the balance),
    (2) the peer-memory halo against the grouped NCCL send/recv: ghost rows and forces bit-identical."""
    import torch
    import torch.distributed as dist

    t = 3
    rig.step(t)
    torch.cuda.synchronize()
    st, halo = rig.st, rig.halo
    sums = torch.zeros(4, dtype=torch.float64, device=rig.dev)
    for f in rig.forces:
        F = f._force[: st.N, :3].double()
        sums[:3] += F.sum(dim=0)
        sums[3] += F.abs().sum()
    dist.all_reduce(sums)
    total, scale = float(sums[:3].abs().max().item()), float(sums[3].item())
    out = {"sum_f_over_sum_abs_f": total / max(scale, 1e-300), "third_law_ok": bool(total <= 1e-6 * scale), "transport": halo.transport}
    ok = out["third_law_ok"]
    if halo.transport == "peer":
        ref = [f._force.clone() for f in rig.forces]
        ghosts = [getattr(st, name)[st.N:].clone() for name in halo.fields]
        halo.transport = "nccl"
        for name in halo.fields:
            getattr(st, name)[st.N:].zero_()
        rig.step(t)
        torch.cuda.synchronize()
        halo.transport = "peer"
        same = all(torch.equal(a, f._force) for a, f in zip(ref, rig.forces)) and all(
            torch.equal(g, getattr(st, name)[st.N:]) for g, name in zip(ghosts, halo.fields))
        flag = torch.tensor([1 if same else 0], device=rig.dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        out["peer_equals_nccl_bitwise"] = bool(flag.item())
        ok = ok and out["peer_equals_nccl_bitwise"]
        rig.step(t)  # leaves the epochs of the peer path consistent on every rank
        torch.cuda.synchronize()
    halo.check()
    out["ok"] = bool(ok)
    return out


def amortised_record(rig, flush, kern_total_ms):
    """What a running simulation pays per step: force kernels + (neighbour list + tile plans) / rebuild period. The rebuild
    is timed on this workload; the period is MEASURED by integrating the workload with md.Integrator (NVE velocity Verlet
    with the workload's own forces, HOOMD's rebuild criterion: any particle moved buffer / 2)."""
    import torch

    from azplugins_b200 import md

    st, nl = rig.st, rig.nl
    ts_list, ts_plan = [], []
    for _ in range(3):
        nl.invalidate()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        nl.compute(st)
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        for f in rig.forces:
            if hasattr(f, "_current_plan"):
                f._current_plan()
        torch.cuda.synchronize()
        ts_list.append(t1 - t0)
        ts_plan.append(time.perf_counter() - t1)
    nlist_ms, plan_ms = min(ts_list) * 1e3, min(ts_plan) * 1e3
    # rebuild period and the whole NVE step: positions are restored afterwards
    keep = (st.pos.clone(), st.vel.clone())
    dt = 0.01 if any(hasattr(f, "dt") for f in rig.forces) else 0.005
    ig = md.Integrator(dt=dt, methods=[md.ConstantVolume()], forces=rig.forces).attach(st)
    ig.run(5)
    torch.cuda.synchronize()
    r0, n_steps = ig.nlist_rebuilds, 40
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    ig.run(n_steps)
    e1.record()
    torch.cuda.synchronize()
    md_ms = e0.elapsed_time(e1) / n_steps
    rebuilds = ig.nlist_rebuilds - r0
    period = n_steps / max(rebuilds, 1)
    st.pos.copy_(keep[0])
    st.vel.copy_(keep[1])
    nl.invalidate()
    nl.compute(st)
    for f in rig.forces:
        if hasattr(f, "_current_plan"):
            f._current_plan()
    torch.cuda.synchronize()
    per_step = kern_total_ms + (nlist_ms + plan_ms) / period
    return {"value": st.N / (per_step * 1e-3), "unit": UNIT, "ms_per_step": per_step, "force_ms": kern_total_ms, "nlist_rebuild_ms": nlist_ms,
            "tile_plan_ms": plan_ms, "rebuild_period_steps": period,
            "rebuild_period_source": f"md.Integrator NVE dt={dt}, {n_steps} steps from the workload's start (velocities N(0, kT=1)), {rebuilds} rebuilds",
            "md_step_ms": md_ms, "md_value": st.N / (md_ms * 1e-3),
            "note": "value = N / (force kernels + (list + tile plans) / period); md_value = whole velocity-Verlet step through md.Integrator "
                    "(integrator kernels, displacement check and rebuilds included, no L2 flush)"}


def run_b200(args):
    import torch
    import torch.distributed as dist

    from azplugins_b200 import _abi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device; the force path has no CPU fallback")
    numa_note = bind_to_gpu_numa_node(local_rank)
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    _abi.load()

    n_per_gpu = args.n
    strong = args.strong and world > 1
    if strong:
        n_per_gpu = max((args.n or DEFAULT_N[args.workload]) // world, 1000)
    rig = Rig(args, args.workload, args.bits, dev, world, rank, n_per_gpu, strong)
    st, forces, halo = rig.st, rig.forces, rig.halo
    flush = None if args.no_flush else torch.empty(384 * 1024 * 1024, dtype=torch.uint8, device=dev)
    W = max(args.warmup, 3)
    K = args.steps
    for t in range(W):
        rig.step(t)
    rig.barrier()
    check = mgpu_check(rig) if world > 1 else None
    launches0 = rig.launches()

    # ---- timed region ----------------------------------------------------------------------------------------------------
    sampler = ClockSampler(local_rank)
    sampler.start()
    t_begin = sampler.mark()
    step_ms, kern_ms = rig.time_steps(K, W, flush, min_seconds=args.min_seconds)
    t_end = sampler.mark()
    clocks = sampler.stop(t_begin, t_end)
    blocks = step_ms.shape[0]
    total_ms = float(step_ms.sum()) / blocks  # mean K-step block
    if world > 1:
        tt = torch.tensor([total_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        total_ms = float(tt.item())
        nn = torch.tensor([st.N], dtype=torch.int64, device=dev)
        dist.all_reduce(nn)
        n_total = int(nn.item())
    else:
        n_total = st.N
    launches = (rig.launches() - launches0) // blocks
    value = n_total * K / (total_ms * 1e-3)

    # ---- roofline of the dominant kernel (forces[0]) -------------------------------------------------------------------------
    peak, peak_src = hbm_peak()
    recs = rig.kernel_records(kern_ms, peak)
    dom = max(recs, key=lambda r: r["kernel_ms"])  # the dominant kernel of the step = the slowest launch
    roofline = {"bound": "hbm", "achieved": dom["achieved"], "peak": peak, "unit": "GB/s", "frac": dom["frac"], "traffic": None,
                "kernel": dom["kernel"], "bytes_per_particle": dom["bytes_per_particle"], "mean_neighbors": rig.nbar, "kernel_ms": dom["kernel_ms"],
                "peak_source": peak_src, "pair_visits_per_s": rig.nbar * st.N / (dom["kernel_ms"] * 1e-3), "kernels": recs,
                "step_bytes_per_particle": sum(r["bytes_per_particle"] for r in recs),
                "step_frac": st.N * sum(r["bytes_per_particle"] for r in recs) / (sum(r["kernel_ms"] for r in recs) * 1e-3) / 1e9 / peak}
    roofs_from_profile(args.workload, args.bits, roofline, dom["kernel"])

    plan = forces[0]._plan
    tile_note = ("gathered from global memory" if plan is None else
                 f"staged per tile in shared memory (azb_tile_plan: {plan.tile_size} particles/tile, <= {plan.max_uniq} distinct neighbours, "
                 f"built in {plan.build_ms:.2f} ms per neighbour-list rebuild, outside the timed region like the list itself; see `amortised`)")
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "strong" if strong else "weak", "vs_baseline": None,
        "dtype": "f32" if args.bits == 32 else "f64", "data": "synthetic",
        "config": shared_config(rig.label, st.N if not strong else DEFAULT_N[args.workload] // world, rig.virial, world, args.bits, strong),
        "clocks": clocks, "roofline": roofline, "gpu_launches": int(launches), "timed_blocks": int(blocks),
        "neighbor_data": tile_note, "mean_neighbors": round(rig.nbar, 2), "threads_per_particle": args.tpp or "auto",
        "nlist_build_s": rig.nlist_build_s, "host_affinity": numa_note,
    }
    if strong:
        line["config"]["particles_per_gpu"] = int(st.N)
    if halo is not None:
        line["halo"] = {"transport": ("peer memory: one push kernel storing into the neighbours' arrays over NVLink, flag-synchronised"
                                      if halo.transport == "peer" else "NCCL send/recv"),
                        "overlap": "ghost-free tiles run while the halo travels", "bytes_sent_per_rank_per_step": halo.bytes_per_update(),
                        "ghosts_rank0": st.n_ghost, "decomposition": "static for the timed steps (no migration inside the force benchmark)"}
        line["mgpu_check"] = check

    # ---- e2e: the same metric through the public API with HOST buffers; every step copies this step's positions
    # (+ velocities / orientations) from pinned host memory and reads the forces (+ torques) back into pinned host memory.
    # Neighbour list, tile plan, tags and parameter tables stay resident between list rebuilds, as they do in HOOMD.
    from azplugins_b200.hostio import HostStep

    tiled = [f for f in forces if getattr(f, "_plan", None) is not None]
    plain = [f for f in forces if f not in tiled]
    pipelined = args.e2e_chunks > 1 and len(tiled) >= 1
    taper = () if args.e2e_taper in ("none", "") else tuple(float(x) / 100.0 for x in args.e2e_taper.split(","))
    hs = HostStep(tiled if pipelined else forces[:1], chunks=args.e2e_chunks, halo=halo, taper=taper)
    rest = plain if pipelined else forces[1:]
    extra_out = [f._force for f in rest]
    pinned_in = hs.pinned_inputs()
    pinned_out = hs.pinned_outputs()
    pinned_extra = [torch.empty(t.shape, dtype=t.dtype, pin_memory=True) for t in extra_out]
    h2d = hs.h2d_bytes
    d2h = hs.d2h_bytes + sum(t.numel() * t.element_size() for t in extra_out)

    def e2e_step(t):
        if pipelined:
            hs.run(t, pinned_in, pinned_out)
            if rest:
                for f in rest:
                    f.compute(t)
                for dst, src in zip(pinned_extra, extra_out):
                    dst.copy_(src, non_blocking=True)
                torch.cuda.current_stream().synchronize()
            return
        for dst, src in zip(hs.inputs, pinned_in):
            dst.copy_(src, non_blocking=True)
        rig.step(t)
        for dst, src in zip(pinned_out + pinned_extra, hs.outputs + extra_out):
            dst.copy_(src, non_blocking=True)
        torch.cuda.current_stream().synchronize()  # the host consumes the forces every step

    for t in range(3):
        e2e_step(t)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    rig.barrier()
    e0.record()
    for k in range(K):
        e2e_step(W + k)
    e1.record()
    rig.barrier()
    e2e_ms = e0.elapsed_time(e1)
    if world > 1:
        tt = torch.tensor([e2e_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        e2e_ms = float(tt.item())
    if pipelined:
        # the pipelined step must deliver the same forces as the plain launches
        rig.step(W + K - 1)
        torch.cuda.synchronize()
        same = all(torch.equal(o.cpu(), p) for o, p in zip(hs.outputs, pinned_out))
        if not same:
            raise RuntimeError("pipelined host step and plain launch disagree")
    line["e2e"] = {"value": n_total * K / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                   "ms_per_step": e2e_ms / K,
                   "note": (f"pipelined in {args.e2e_chunks} chunks (the last ones hold {args.e2e_taper} % of the rows): pinned host pos(+vel/orientation) -> device, tile-range launches of "
                            f"{len(tiled)} force object(s) as inputs arrive"
                            + (", ghost exchange behind the last input chunk, then the tiles that read ghosts" if halo is not None else "")
                            + ", force(+torque) arrays -> pinned host; checked bit-identical to the plain launches"
                            if pipelined else
                            "pinned host pos(+vel/orientation) -> device, (ghost exchange,) kernel(s), force(+torque) -> pinned host, every step")}

    extras = rank == 0 and world == 1 and not args.no_extras
    if extras and flush is not None:
        # back-to-back (no L2 flush) figure for context
        torch.cuda.synchronize()
        e0.record()
        for k in range(K):
            rig.step(k)
        e1.record()
        torch.cuda.synchronize()
        line["value_back_to_back_no_flush"] = st.N * K / (e0.elapsed_time(e1) * 1e-3)
    if extras:
        try:
            line["amortised"] = amortised_record(rig, flush, float(sum(r["kernel_ms"] for r in recs)))
        except Exception as exc:  # noqa: BLE001 - a sub-record never costs the headline
            line["amortised"] = {"value": None, "error": f"{type(exc).__name__}: {exc}"}

    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        nl = rig.nl
        nl_h = (nl.n_neigh.cpu().numpy().view(np.uint32), nl.head_list.cpu().numpy().view(np.uint64), nl.nlist.cpu().numpy().view(np.uint32))
        try:
            v, cores, kind, sec, reps = time_reference_cpu(args.workload, rig.snap, args.bits, nl_h, args.cpu_seconds)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": kind,
                                    "sample": f"{reps} evaluations of the whole workload ({st.N} particles, same neighbour list), "
                                              f"{sec * 1e3:.1f} ms each, OpenMP over particles, Scalar={'float' if args.bits == 32 else 'double'}"}
        except Exception as exc:  # the baseline is a report, never a reason to lose the GPU number
            line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "port", "sample": f"failed: {exc}"}
        if extras:
            try:
                line["cpu_baseline"]["one_core"] = one_core_baseline(args.workload, args.cpu_seconds)
            except Exception as exc:  # noqa: BLE001
                line["cpu_baseline"]["one_core"] = {"value": None, "error": f"{type(exc).__name__}: {exc}"}
        del nl_h

    if halo is not None:
        halo.check()  # a flag wait that timed out would have left stale ghosts

    # ---- sub-records (N = 1): double precision, the other configurations; (N > 1): strong scaling of the same workload -------
    del hs, pinned_in, pinned_out, pinned_extra
    if extras:
        if args.bits == 32:
            line["f64"] = sub_record(args, args.workload, 64, dev, flush, peak, steps=min(K, 5), n=args.n)
        line["others"] = []
        del rig
        torch.cuda.empty_cache()
        for w in ("dpd", "plj4m", "melt", "patchy"):
            if w == args.workload:
                continue
            try:
                line["others"].append(sub_record(args, w, args.bits, dev, flush, peak, steps=min(K, 5)))
            except Exception as exc:  # noqa: BLE001
                line["others"].append({"workload": w, "error": f"{type(exc).__name__}: {exc}"})
            torch.cuda.empty_cache()
    elif world > 1 and not strong and not args.no_extras:
        del rig
        torch.cuda.empty_cache()
        try:
            line["strong"] = strong_record(args, dev, world, rank, flush, K, W)
        except Exception as exc:  # noqa: BLE001
            line["strong"] = {"value": None, "error": f"{type(exc).__name__}: {exc}"}
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def sub_record(args, workload, bits, dev, flush, peak, steps, n=0):
    """A short device-timed measurement of another workload / precision on this GPU (same timing rules, fewer steps)."""
    import torch

    rig = Rig(args, workload, bits, dev, 1, 0, n)
    for t in range(3):
        rig.step(t)
    torch.cuda.synchronize()
    step_ms, kern_ms = rig.time_steps(steps, 3, flush)
    recs = rig.kernel_records(kern_ms, peak)
    ms = float(step_ms.mean())
    out = {"workload": rig.label, "dtype": "f32" if bits == 32 else "f64", "particles": rig.st.N, "mean_neighbors": round(rig.nbar, 2), "steps": steps,
           "ms_per_step": ms, "value": rig.st.N / (ms * 1e-3), "unit": UNIT, "kernels": recs, "kernel_ms": recs[0]["kernel_ms"], "frac": recs[0]["frac"],
           "step_frac": rig.st.N * sum(r["bytes_per_particle"] for r in recs) / (sum(r["kernel_ms"] for r in recs) * 1e-3) / 1e9 / peak}
    for f in rig.forces:
        if hasattr(f, "check_flags"):
            f.check_flags()
    return out


def strong_record(args, dev, world, rank, flush, K, W):
    """Strong scaling of the same workload: its single-GPU particle count split over the ranks (second record of the line, so
    that the driver's scaling run captures both)."""
    import torch
    import torch.distributed as dist

    total = args.n or DEFAULT_N[args.workload]
    rig = Rig(args, args.workload, args.bits, dev, world, rank, max(total // world, 1000), True)
    for t in range(W):
        rig.step(t)
    rig.barrier()
    check = mgpu_check(rig)
    step_ms, _ = rig.time_steps(K, W, flush)
    tt = torch.tensor([float(step_ms.sum()) / step_ms.shape[0]], dtype=torch.float64, device=dev)
    dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    nn = torch.tensor([rig.st.N], dtype=torch.int64, device=dev)
    dist.all_reduce(nn)
    ms = float(tt.item())
    return {"scaling": "strong", "particles_total": int(nn.item()), "particles_per_gpu": rig.st.N, "ms_per_step": ms / K,
            "value": int(nn.item()) * K / (ms * 1e-3), "unit": UNIT, "mgpu_check": check}


def one_core_baseline(workload, seconds):
    """What ONE HOOMD CPU rank executes: Scalar = double (HOOMD's default), half neighbour list + Newton's third
    law, one thread — on a bounded sample: the same recipe at <= 250 000 particles (the per-particle cost does not depend on
    the system size beyond cache effects, which favour the small sample)."""
    n = min(DEFAULT_N[workload], 250_000)
    label, snap, _, buffer = make_workload(workload, n)
    rc = list_cutoff(workload, snap)
    nl = oracle_nlist(snap, 64, rc + buffer, bond_exclusions(snap))
    v, cores, kind, sec, reps = time_reference_cpu(workload, snap, 64, nl, min(seconds, 8.0), nthreads=1, half_list=True)
    return {"value": v, "unit": UNIT, "cores": 1, "kind": kind,
            "sample": f"{reps} evaluations of the same recipe at N={snap.N}, {sec * 1e3:.0f} ms each, Scalar=double, half list + third law, 1 thread"}


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
