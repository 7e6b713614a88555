"""The ctypes mirrors in azplugins_b200/_abi.py against include/azb200.h as a C compiler lays it out: size of every struct
and offset of every field. (A drifted mirror does not fail loudly — it hands the kernels garbage.) CPU only: gcc compiles a
probe that includes the header and prints sizeof / offsetof."""

import ctypes as C
import os
import shutil
import subprocess

import pytest

from azplugins_b200 import _abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

MIRRORS = {
    "azb_box": _abi.Box,
    "azb_tile_plan": _abi.TilePlan,
    "azb_tile_build_args": _abi.TileBuildArgs,
    "azb_pair_args": _abi.PairArgs,
    "azb_dpd_args": _abi.DPDArgs,
    "azb_aniso_args": _abi.AnisoArgs,
    "azb_bond_args": _abi.BondArgs,
    "azb_nlist_view": _abi.NlistView,
    "azb_nlist_plan_status": _abi.NlistPlanStatus,
    "azb_nlist_args": _abi.NlistArgs,
    "azb_nve_args": _abi.NveArgs,
    "azb_barrier_args": _abi.BarrierArgs,
    "azb_halo_route": _abi.HaloRoute,
    "azb_halo_push_args": _abi.HaloPushArgs,
    "azb_host_buffer": _abi.HostBuffer,
    "azb_host_launch": _abi.HostLaunch,
    "azb_host_out_range": _abi.HostOutRange,
    "azb_host_force": _abi.HostForce,
    "azb_host_step_args": _abi.HostStepArgs,
}


@pytest.mark.skipif(shutil.which("gcc") is None, reason="needs gcc")
def test_ctypes_mirrors_match_the_header(tmp_path):
    lines = ["#include <stddef.h>", "#include <stdio.h>", '#include "azb200.h"', "int main(void) {"]
    for cname, cls in MIRRORS.items():
        lines.append(f'  printf("{cname} . %zu\\n", sizeof({cname}));')
        for field in cls._fields_:
            lines.append(f'  printf("{cname} {field[0]} %zu\\n", offsetof({cname}, {field[0]}));')
    lines += ["  return 0;", "}"]
    src = tmp_path / "probe.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "probe"
    subprocess.run(["gcc", "-std=c11", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout
    seen = 0
    for row in out.splitlines():
        cname, field, value = row.split()
        cls = MIRRORS[cname]
        want = C.sizeof(cls) if field == "." else getattr(cls, field).offset
        assert int(value) == want, (cname, field, int(value), want)
        seen += 1
    assert seen == sum(1 + len(c._fields_) for c in MIRRORS.values())
