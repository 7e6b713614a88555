"""Compile-only guard of hoomd_glue/ (SURVEY.md §8 f3): the kernel-driver stubs that replace the reference's four
*GPUKernel.cu.inc, the tiled host class and its exporter are stamped per evaluator exactly as CMake's configure_file() would,
compiled against hoomd_glue/mock (HOOMD v5's argument structs as recalled) together with the reference's UNMODIFIED evaluator
headers, and linked with libazb200.so with no undefined symbol left. CPU only; needs /root/reference (skipped on the GPU box).
A build against a real HOOMD install stays unverified here: HOOMD is not installable on this box."""

import os
import shutil
import subprocess
import sysconfig

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GLUE = os.path.join(ROOT, "hoomd_glue")
REF_SRC = "/root/reference/src"
LIB = os.path.join(ROOT, "azplugins_b200", "libazb200.so")

UNITS = [  # (template, evaluators, symbol fragment that must be defined in the object)
    ("PotentialPairGPUKernel.cu.inc", ["Colloid", "ExpandedYukawa", "Hertz", "PerturbedLennardJones"], "gpu_compute_pair_forces"),
    ("PotentialPairDPDThermoGPUKernel.cu.inc", ["GeneralWeight"], "gpu_compute_dpd_forces"),
    ("AnisoPotentialPairGPUKernel.cu.inc", ["TwoPatchMorse"], "gpu_compute_pair_aniso_forces"),
    ("PotentialBondGPUKernel.cu.inc", ["DoubleWell", "Quartic"], "gpu_compute_bond_forces"),
    ("export_PotentialPairGPU.cc.inc", ["Colloid", "ExpandedYukawa", "Hertz", "PerturbedLennardJones"], "export_PotentialPair"),
]

pytestmark = pytest.mark.skipif(not os.path.isdir(REF_SRC) or shutil.which("g++") is None or not os.path.exists(LIB),
                                reason="needs /root/reference, g++ and the built libazb200.so")


def _flags():
    import pybind11

    return ["-std=c++17", "-O0", "-fPIC", "-Wall", "-Wno-unused-variable", "-DENABLE_HIP", f"-I{os.path.join(GLUE, 'mock')}", f"-I{GLUE}", f"-I{REF_SRC}",
            f"-I{os.path.join(ROOT, 'include')}", f"-I{pybind11.get_include()}", f"-I{sysconfig.get_paths()['include']}"]


def test_glue_units_compile_against_mock_and_link_with_the_library(tmp_path):
    objs = []
    for template, evaluators, symbol in UNITS:
        text = open(os.path.join(GLUE, template)).read()
        assert "@_evaluator@" in text
        for ev in evaluators:
            src = tmp_path / (template.replace(".inc", "").replace(".cu", f"_{ev}.cc").replace("GPU.cc", f"GPU_{ev}.cc"))
            src.write_text(text.replace("@_evaluator@", ev))
            obj = str(src) + ".o"
            r = subprocess.run(["g++", *_flags(), "-x", "c++", "-c", str(src), "-o", obj], capture_output=True, text=True)
            assert r.returncode == 0, f"{template} [{ev}]\n{r.stderr[-3000:]}"
            names = subprocess.run(["nm", "-C", "--defined-only", obj], capture_output=True, text=True).stdout
            assert symbol in names and ev in names, (template, ev)
            objs.append(obj)
    # module.cc: the exporters it declares are the ones the stamped units define (GPU pair) or the reference defines (rest)
    r = subprocess.run(["g++", *_flags(), "-c", os.path.join(GLUE, "module.cc"), "-o", str(tmp_path / "module.o")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    # link: every azb_* call of the stubs must resolve in libazb200.so; nothing else may be missing except Python itself
    out = str(tmp_path / "libglue_check.so")
    r = subprocess.run(["g++", "-shared", "-o", out, *objs, f"-L{os.path.dirname(LIB)}", "-lazb200", f"-Wl,-rpath,{os.path.dirname(LIB)}",
                        "-Wl,--unresolved-symbols=report-all", "-Wl,--warn-unresolved-symbols"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    unresolved = [ln for ln in r.stderr.splitlines() if "undefined reference" in ln and "azb_" in ln]
    assert not unresolved, unresolved
    und = subprocess.run(["nm", "-D", "--undefined-only", out], capture_output=True, text=True).stdout
    wanted = {"azb_pair_forces", "azb_dpd_forces", "azb_aniso_forces", "azb_bond_forces", "azb_tile_cache_get", "azb_tile_cache_invalidate"}
    assert wanted <= {ln.split()[-1] for ln in und.splitlines() if ln.strip()}, und
    exported = subprocess.run(["nm", "-D", "--defined-only", LIB], capture_output=True, text=True).stdout
    for name in wanted:
        assert f" T {name}" in exported, name


def test_cmake_lists_name_every_template():
    text = open(os.path.join(GLUE, "CMakeLists.txt")).read()
    for template, _, _ in UNITS:
        assert template in text
    assert "compute_100a" in text and "libazb200.so" in text
