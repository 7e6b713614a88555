"""Multi-GPU parity through torchrun (skipped when fewer than 2 GPUs are visible)."""

import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu


def _ngpu():
    import torch

    return torch.cuda.device_count()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_decomposed_dpd_forces_match_oracle(world):
    if _ngpu() < world:
        pytest.skip(f"needs {world} GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(29540 + world), os.path.join(root, "tests", "mgpu_force_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "mgpu force check ok" in r.stdout


@pytest.mark.parametrize("world", [2, 4, 8])
def test_decomposed_md_with_migration_matches_one_gpu(world):
    """md.Integrator on 2 / 4 / 8 GPUs (migration + ghost re-selection + ghost bonds on every list rebuild) against the 1-GPU
    trajectory: C1 for 1000 steps, a bead-spring melt with both bond potentials for 300."""
    if _ngpu() < world:
        pytest.skip(f"needs {world} GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(29560 + world), os.path.join(root, "tests", "mgpu_md_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count("mgpu md check ok") == 2
