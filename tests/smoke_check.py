"""One small invocation of each kernel family on cuda:0, checked against the CPU oracle (used by smoke())."""

import numpy as np

import gpu_cases as GC


def run(verbose=False):
    import torch

    res = {}
    par = {(0, 0): dict(A=25.0, gamma=4.5, s=2.0)}
    f, hs, _ = GC.run_dpd(par, {(0, 0): 1.0}, 32, n=4000, ntypes=1, timestep=3)
    res["dpd_f32"] = GC.error_metrics(f.forces, hs.force[:, :3])[0]
    f, hs = GC.run_pair("PerturbedLennardJones", {(0, 0): dict(epsilon=1.0, sigma=1.0, attraction_scale_factor=0.5)}, {(0, 0): 3.0}, 32,
                        n=4000, rho=0.8, ntypes=1, buffer=0.4)
    res["plj_f32"] = GC.error_metrics(f.forces, hs.force[:, :3])[0]
    f, hs = GC.run_bond("Quartic", [dict(k=1434.3, r_0=1.5, b_1=-0.7589, b_2=0.0, U_0=67.2234, sigma=1.0, epsilon=1.0, delta=0.0)], 32, n=4000)
    res["quartic_f32"] = GC.error_metrics(f.forces, hs.force[:, :3])[0]
    torch.cuda.synchronize()
    if verbose:
        print("smoke: max normalised force error vs oracle:", res)
    for k, v in res.items():
        assert np.isfinite(v) and v < 1e-4, (k, v)
    return res
