"""N > 1 host logic on CPU: world_size 2 / 4 / 8 gloo processes decompose one seeded periodic system, exchange ghosts and
check (by tag) that every local particle sees exactly the neighbours it has in the global system, that ghost
types / velocities / tags travel, and that the per-step halo update moves ghosts with their owners."""

import socket

import numpy as np
import pytest
import torch.multiprocessing as mp

import domain_worker
from azplugins_b200 import Box
from azplugins_b200.domain import Decomposition


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("world", [2, 4, 8])
def test_halo_exchange_gloo(world):
    mp.spawn(domain_worker.run, args=(world, _free_port()), nprocs=world, join=True)


def test_decomposition_geometry():
    box = Box(20.0, 10.0, 10.0)
    for world, grid in ((1, (1, 1, 1)), (2, (2, 1, 1)), (4, (2, 2, 1)), (8, (2, 2, 2))):
        decs = [Decomposition(world, r, box) for r in range(world)]
        assert decs[0].grid == grid
        rng = np.random.default_rng(1)
        pos = rng.uniform(-0.5, 0.5, size=(2000, 3)) * box.L
        owners = np.sum([d.owns(pos) for d in decs], axis=0)
        assert np.all(owners == 1)
        for d in decs:
            assert d.rank_of(d.coords) == d.rank
            for dr in d.active_directions():
                peer, shift = d.neighbor(dr)
                back, shift_back = decs[peer].neighbor(tuple(-x for x in dr))
                assert back == d.rank and np.allclose(shift, -shift_back)
    with pytest.raises(ValueError):
        Decomposition(3, 0, box)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_migration_ghost_reselection_and_ghost_bonds_gloo(world):
    """On a list rebuild: leavers migrate (variable-size messages, sizes first), ghosts are re-selected, bonds are re-indexed
    with ghost partners — four rebuilds with every particle displaced by up to 0.19 each time."""
    mp.spawn(domain_worker.run_migration, args=(world, _free_port()), nprocs=world, join=True)
