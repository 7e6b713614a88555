"""bench.py — force-compute throughput of the B200 engine (and, with --impl reference, of the reference's CPU path).

One "step" = one evaluation of the hot path (the force kernel launch through the C-ABI) over the whole synthetic
configuration. Default workload: BASELINE.json configs[1] — DPDGeneralWeight fluid, ~1M particles, rho = 3,
r_cut = 1.0, buffer 0.4, Philox random forces — Scalar = float. Prints ONE JSON line (rank 0).

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


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="dpd", choices=["dpd", "plj", "plj4m", "patchy", "hertz", "colloid", "melt", "eydpd"])
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
    ap.add_argument("--no-extras", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    ap.add_argument("--e2e-chunks", type=int, default=8, help="chunks of the pipelined host step (1 = copy, compute, copy)")
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
            return f"process bound to NUMA node {node} (cpus {cpus}) of GPU {bus}"
        return f"GPU {bus}: no usable local cpus ({cpus})"
    except Exception as exc:  # noqa: BLE001 - affinity is an optimisation, never a failure
        return f"not bound ({type(exc).__name__})"


# ---------------------------------------------------------------------------------------------------------------
def make_workload(name, n_override, rank=0):
    """Returns (label, snapshot, setup) where setup(nlist) -> list of force objects (first = dominant kernel)."""
    from azplugins_b200 import bond, pair, workloads as W

    if name == "dpd":
        n = n_override or 1_000_188
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
            q = bond.Quartic()
            q.params["quartic"] = W.QUARTIC_PARAMS
            q.params["doublewell"] = dict(W.QUARTIC_PARAMS)  # Quartic object sees both bond types; see DESIGN.md
            return [f, q]

        return f"C3 bead-spring melt N={snap.N} (chains of 100) PerturbedLJ r_cut=3.0 + bond.Quartic", snap, setup, 0.4
    if name == "eydpd":
        n = n_override or 8_000_000
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
                "(roofline: the DPD kernel)"), snap, setup, 0.4
    raise ValueError(name)


def algorithmic_bytes_per_particle(force, nbar, bits, virial):
    """SURVEY.md §8 d: compulsory HBM bytes per particle-step of one kernel."""
    from azplugins_b200 import _abi

    s = bits // 8
    S4 = 4 * s
    v = 6 * s if virial else 0
    if force._kind == _abi.KIND_BOND:
        return S4 + 4 + S4 + v + 12 * nbar  # nbar = bond-table slots per particle
    b = S4 + 4 + 8 + S4 + v + 4 * nbar
    if force._kind == _abi.KIND_DPD:
        b += S4 + 4
    if force._kind == _abi.KIND_ANISO:
        b += 2 * S4
    return b


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
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
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


def time_reference_cpu(workload, snap, bits, nl_arrays, seconds, steps=None, warmup=1):
    """The reference's own CPU implementation of the path: its evaluator headers compiled unmodified (oracle/_ref) —
    or the C port when /root/reference was not available at build time — inside the restated HOOMD host loop, full
    neighbour list, OpenMP over particles on every host core. Returns (particle-steps/s, cores, kind, seconds/step)."""
    import __graft_entry__ as g

    g.build(only_if_missing=True)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import cases
    import oracle_api as O
    from azplugins_b200 import workloads as W

    hs, _ = host_arrays_for_oracle(snap, bits, *nl_arrays)
    which = "ref" if O.have_ref() else "port"
    os.sched_setaffinity(0, ALL_CPUS)  # undo the GPU-local binding of the copy path: the reference gets all host cores
    cores = len(ALL_CPUS)
    T = len(snap.types)
    names = snap.types

    def tables(spec, pdict, rc):
        par_by = {(min(i, j), max(i, j)): pdict[tuple(sorted((names[i], names[j])))] for i in range(T) for j in range(T)}
        rc_by = {(min(i, j), max(i, j)): (rc[tuple(sorted((names[i], names[j])))] if isinstance(rc, dict) else rc) for i in range(T) for j in range(T)}
        return cases.pair_tables(spec, par_by, rc_by, T, bits)

    if workload == "dpd":
        par, rcsq, _, _ = tables(cases.SPEC_BY_API["DPDGeneralWeight"], W.DPD_PARAMS, 1.0)
        step = lambda t: hs.dpd(par, rcsq, seed=42, timestep=t, dt=0.01, kT=1.0, nthreads=cores, which=which)  # noqa: E731
    elif workload in ("plj", "plj4m", "melt"):
        spec = cases.SPEC_BY_API["PerturbedLennardJones"]
        par, rcsq, ronsq, _ = tables(spec, W.PLJ_PARAMS, 3.0)
        step = lambda t: hs.pair(spec.potential_id, par, rcsq, ronsq, nthreads=cores, which=which)  # noqa: E731
    elif workload in ("hertz", "colloid"):
        spec = cases.SPEC_BY_API["Hertz" if workload == "hertz" else "Colloid"]
        pdict, rc = W.hertz_tables() if workload == "hertz" else (W.COLLOID_PARAMS, W.COLLOID_RCUT)
        par, rcsq, ronsq, _ = tables(spec, pdict, rc)
        step = lambda t: hs.pair(spec.potential_id, par, rcsq, ronsq, nthreads=cores, which=which)  # noqa: E731
    elif workload == "patchy":
        par, rcsq, _, _ = tables(cases.SPEC_BY_API["TwoPatchMorse"], W.TPM_PARAMS, 1.6)
        step = lambda t: hs.aniso(par, rcsq, nthreads=cores, which=which)  # noqa: E731
    elif workload == "eydpd":
        par_d, rcsq_d, _, _ = tables(cases.SPEC_BY_API["DPDGeneralWeight"], W.C5_DPD_PARAMS, 1.0)
        spec_y = cases.SPEC_BY_API["ExpandedYukawa"]
        par_y, rcsq_y, ronsq_y, _ = tables(spec_y, W.EY_PARAMS, 2.0)

        def step(t):
            hs.dpd(par_d, rcsq_d, seed=42, timestep=t, dt=0.01, kT=1.0, nthreads=cores, which=which)
            hs.pair(spec_y.potential_id, par_y, rcsq_y, ronsq_y, nthreads=cores, which=which)
    else:
        raise ValueError(workload)
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
        elif k >= 3 and time.perf_counter() - t_begin > seconds:
            break
    best = float(np.mean(times))
    return hs.N / best, cores, ("reference" if which == "ref" else "port"), best, len(times)


def oracle_nlist(snap, bits, r_list):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_api as O

    hs = O.HostSystem(snap.box, snap.position, snap.typeid, len(snap.types), bits)
    hs.build_nlist(r_list)
    return hs.n_neigh, hs.head_list, hs.nlist


def run_reference(args):
    """--impl reference: the reference's CPU path on this box's host cores (rank 0 only)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    label, snap, _, buffer = make_workload(args.workload, args.n)
    if args.workload in ("hertz", "colloid"):
        from azplugins_b200 import workloads as W

        by_pair = W.hertz_tables()[1] if args.workload == "hertz" else W.COLLOID_RCUT
        names = snap.types
        rc = np.array([[by_pair[tuple(sorted((a, b)))] for b in names] for a in names])
    else:
        rc = {"dpd": 1.0, "plj": 3.0, "plj4m": 3.0, "melt": 3.0, "patchy": 1.6, "eydpd": 2.0}[args.workload]
    nl = oracle_nlist(snap, args.bits, rc + buffer)  # input preparation, not timed
    value, cores, kind, sec, k = time_reference_cpu(args.workload, snap, args.bits, nl, 0.0, steps=args.steps, warmup=args.warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32" if args.bits == 32 else "f64", "data": "synthetic",
        "config": {"workload": label, "neighbor_list": "full, HOOMD layout", "host": "CPU only"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind,
                         "sample": f"every step = the whole workload ({snap.N} particles), OpenMP over particles, full neighbour list"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


# ---------------------------------------------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist

    import azplugins_b200 as az
    from azplugins_b200 import _abi, nlist

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
    if args.strong and world > 1:
        # strong scaling: the total is fixed (the workload's default size, or --n), every rank takes 1/world of it
        total = args.n or {"dpd": 1_000_188, "plj": 32_000, "plj4m": 4_000_000, "patchy": 2_048_000, "hertz": 2_048_000, "colloid": 1_000_000, "melt": 4_000_000, "eydpd": 8_000_000}[args.workload]
        n_per_gpu = max(total // world, 1000)
    label, snap, setup, buffer = make_workload(args.workload, n_per_gpu, rank)
    nl = nlist.Cell(buffer=buffer)
    forces = setup(nl)
    halo = None
    if world > 1:
        # weak scaling: every rank generates its own sub-domain (same recipe, rank-dependent seed) of one periodic global
        # box of grid x sub-box; ghosts within r_cut + buffer are exchanged with the neighbouring domains every step
        from azplugins_b200.domain import GRIDS, Decomposition, HaloExchange

        if args.workload == "melt":
            raise SystemExit("melt: bonds across domains are not supported in the multi-GPU bench")
        grid = GRIDS[world]
        sub = snap.box
        gbox = az.Box(sub.Lx * grid[0], sub.Ly * grid[1], sub.Lz * grid[2])
        dec = Decomposition(world, rank, gbox)
        snap.position += (dec.lo + dec.hi) / 2
        snap.tag = (snap.tag.astype(np.uint64) + rank * snap.N).astype(np.uint32)
        snap.box = gbox
    st = az.State(snap, device=dev, scalar_bits=args.bits, seed=42)
    if world > 1:
        kinds = {f._kind for f in forces}
        fields = ("pos",) + (("vel",) if _abi.KIND_DPD in kinds else ()) + (("orientation",) if _abi.KIND_ANISO in kinds else ())
        halo = HaloExchange(dec, st, float(nl.r_list(st.types).max()), fields=fields)
    virial = not args.no_virial
    for f in forces:
        f.compute_virial = virial
        f.threads_per_particle = args.tpp
        f.block_size = args.block
        f.use_tiles = not args.no_tiles
        f.tile_size = args.tile_size
        f._attach(st)
    t0 = time.perf_counter()
    nl.compute(st)
    torch.cuda.synchronize()
    nlist_build_s = time.perf_counter() - t0
    nbar = nl.mean_neighbors(st.N)

    flush = None if args.no_flush else torch.empty(384 * 1024 * 1024, dtype=torch.uint8, device=dev)

    def step(t):
        if halo is not None:
            halo.step(forces[0], t)  # ghost refresh overlapped with the ghost-free tiles (azplugins_b200/domain.py)
            for f in forces[1:]:
                f.compute(t)
            return
        for f in forces:
            f.compute(t)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for t in range(max(args.warmup, 3)):
        step(t)
    barrier()
    launches0 = sum(f.launches for f in forces) + (halo.launches if halo is not None else 0)

    # ---- timed region: K steps, CUDA events on the launching stream around every step; L2 flushed in between ------
    sampler = ClockSampler(local_rank)
    sampler.start()
    K = args.steps
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    kev = [[(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in forces] for _ in range(K)]
    barrier()
    for k in range(K):
        if flush is not None:
            flush.fill_(k & 0xFF)
        ev[k][0].record()
        if halo is not None:
            kev[k][0][0].record()
            halo.step(forces[0], args.warmup + k)
            kev[k][0][1].record()
        for fi, f in enumerate(forces):
            if halo is not None and fi == 0:
                continue
            kev[k][fi][0].record()
            f.compute(args.warmup + k)
            kev[k][fi][1].record()
        ev[k][1].record()
    barrier()
    clocks = sampler.stop()
    step_ms = np.array([a.elapsed_time(b) for a, b in ev])
    kern_ms = np.array([[a.elapsed_time(b) for a, b in row] for row in kev])
    total_ms = float(step_ms.sum())
    if world > 1:
        tt = torch.tensor([total_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        total_ms = float(tt.item())
        nn = torch.tensor([st.N], dtype=torch.int64, device=dev)
        dist.all_reduce(nn)
        n_total = int(nn.item())
    else:
        n_total = st.N
    launches = sum(f.launches for f in forces) + (halo.launches if halo is not None else 0) - launches0
    value = n_total * K / (total_ms * 1e-3)

    # ---- roofline of the dominant kernel (forces[0]) ------------------------------------------------------------------
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (measured copy, burst)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    dom = forces[0]
    Bpp = algorithmic_bytes_per_particle(dom, nbar, args.bits, virial)
    dom_ms = float(kern_ms[:, 0].mean())
    achieved = st.N * Bpp / (dom_ms * 1e-3) / 1e9
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                "kernel": type(dom).__name__, "bytes_per_particle": Bpp, "mean_neighbors": nbar, "kernel_ms": dom_ms, "peak_source": peak_src,
                "pair_visits_per_s": nbar * st.N / (dom_ms * 1e-3)}
    traffic_path = os.path.join(ROOT, "profiles", f"traffic_{args.workload}_f{args.bits}.json")
    if os.path.exists(traffic_path):
        roofline["traffic"] = json.load(open(traffic_path)).get("dram_bytes_per_launch")

    plan = dom._plan
    tile_note = ("gathered from global memory" if plan is None else
                 f"staged per tile in shared memory (azb_tile_plan: {plan.tile_size} particles/tile, <= {plan.max_uniq} distinct neighbours, "
                 f"built in {plan.build_ms:.2f} ms per neighbour-list rebuild, outside the timed region like the list itself)")
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": max(args.warmup, 3),
        "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "strong" if (args.strong and world > 1) else "weak", "vs_baseline": None,
        "dtype": "f32" if args.bits == 32 else "f64", "data": "synthetic",
        "config": {"workload": label, "particles_per_gpu": st.N, "mean_neighbors": round(nbar, 2), "compute_virial": virial,
                   "threads_per_particle": args.tpp or "auto", "neighbor_list": "full, HOOMD layout, built on GPU (azb_nlist_*)",
                   "neighbor_data": tile_note,
                   "l2": "flushed between timed steps (384 MiB fill)" if flush is not None else "not flushed; working set > L2",
                   "parallelism": "1 GPU" if world == 1 else
                   f"{world} GPUs, spatial domains {GRIDS[world]}, ghost halo every step over {'peer memory (one push kernel storing into the neighbours\' arrays over NVLink, flag-synchronised)' if halo.transport == 'peer' else 'NCCL send/recv'}, overlapped with the ghost-free tiles ({halo.bytes_per_update()} B sent/rank/step, "
                   f"{st.n_ghost} ghosts on rank 0)"},
        "clocks": clocks, "roofline": roofline, "gpu_launches": int(launches),
        "nlist_build_s": nlist_build_s, "host_affinity": numa_note,
    }

    # ---- e2e: the same metric through the public API with HOST buffers; every step copies this step's positions
    # (+ velocities / orientations) from pinned host memory and reads the forces (+ torques) back into pinned host memory.
    # Neighbour list, tile plan, tags and parameter tables stay resident between list rebuilds, as they do in HOOMD.
    # 1 GPU, one force: the step is pipelined (azplugins_b200/hostio.py: chunked H2D, tile-range launches as their inputs
    # arrive, D2H of finished chunks). Several forces or several GPUs (ghost exchange needs all local positions first):
    # copy, exchange, compute, copy.
    from azplugins_b200.hostio import HostStep

    hs = HostStep(dom, chunks=args.e2e_chunks)
    extra_out = [f._force for f in forces[1:]]
    pinned_in = hs.pinned_inputs()
    pinned_out = hs.pinned_outputs()
    pinned_extra = [torch.empty(t.shape, dtype=t.dtype, pin_memory=True) for t in extra_out]
    h2d = hs.h2d_bytes
    d2h = hs.d2h_bytes + sum(t.numel() * t.element_size() for t in extra_out)
    pipelined = world == 1 and len(forces) == 1 and args.e2e_chunks > 1

    def e2e_step(t):
        if pipelined:
            hs.run(t, pinned_in, pinned_out)
            return
        for dst, src in zip(hs.inputs, pinned_in):
            dst.copy_(src, non_blocking=True)
        step(t)
        for dst, src in zip(pinned_out + pinned_extra, hs.outputs + extra_out):
            dst.copy_(src, non_blocking=True)
        torch.cuda.current_stream().synchronize()  # the host consumes the forces every step

    for t in range(3):
        e2e_step(t)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for k in range(K):
        e2e_step(args.warmup + k)
    e1.record()
    barrier()
    e2e_ms = e0.elapsed_time(e1)
    if world > 1:
        tt = torch.tensor([e2e_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        e2e_ms = float(tt.item())
    if pipelined:
        # the pipelined step must deliver the same forces as the plain launch
        step(args.warmup + K - 1)
        torch.cuda.synchronize()
        same = all(torch.equal(o.cpu(), p) for o, p in zip(hs.outputs, pinned_out))
        if not same:
            raise RuntimeError("pipelined host step and plain launch disagree")
    line["e2e"] = {"value": n_total * K / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                   "ms_per_step": e2e_ms / K,
                   "note": (f"pipelined in {args.e2e_chunks} chunks: pinned host pos(+vel/orientation) -> device, tile-range launches as inputs arrive, "
                            "force(+torque) -> pinned host" if pipelined else
                            "pinned host pos(+vel/orientation) -> device, (ghost exchange,) kernel(s), force(+torque) -> pinned host, every step")}

    if rank == 0 and world == 1:
        if not args.no_extras and flush is not None:
            # back-to-back (no L2 flush) figure for context
            torch.cuda.synchronize()
            e0.record()
            for k in range(K):
                step(k)
            e1.record()
            torch.cuda.synchronize()
            line["value_back_to_back_no_flush"] = st.N * K / (e0.elapsed_time(e1) * 1e-3)

        if not args.no_cpu_baseline:
            nn_h = nl.n_neigh.cpu().numpy().view(np.uint32)
            hd_h = nl.head_list.cpu().numpy().view(np.uint64)
            nl_h = nl.nlist.cpu().numpy().view(np.uint32)
            try:
                v, cores, kind, sec, reps = time_reference_cpu(args.workload, snap, args.bits, (nn_h, hd_h, nl_h), args.cpu_seconds)
                line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": kind,
                                        "sample": f"{reps} evaluations of the whole workload ({st.N} particles, same neighbour list), "
                                                  f"{sec * 1e3:.1f} ms each, OpenMP over particles, Scalar={'float' if args.bits == 32 else 'double'}"}
            except Exception as exc:  # the baseline is a report, never a reason to lose the GPU number
                line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "port", "sample": f"failed: {exc}"}
    if halo is not None:
        halo.check()  # a flag wait that timed out would have left stale ghosts
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
