"""Particle data in HOOMD-blue's GPU array layouts.

HOOMD is not installable here, so this module carries the minimum of ``hoomd.Snapshot`` / ``hoomd.State`` that the
force path reads: ``Scalar4`` position+type, velocity+mass, orientation quaternions, tags, the box, and the bond
table in the per-particle GPU form (``group_storage<2>`` at ``[slot * pitch + i]``) that HOOMD's ``BondData``
hands to ``gpu_compute_bond_forces`` (/root/reference/src/PotentialBondGPUKernel.cu.inc:25-29).
Arrays are torch tensors so that they can live in B200 HBM; torch is plumbing only (allocation, copies, streams).
"""

import numpy as np
import torch

from . import _abi


def scalar_dtype(bits):
    if bits == 32:
        return np.float32, np.int32, torch.float32
    if bits == 64:
        return np.float64, np.int64, torch.float64
    raise ValueError("scalar_bits must be 32 or 64")


class Box:
    """Periodic triclinic box, centred on the origin (``hoomd.Box``)."""

    def __init__(self, Lx, Ly=None, Lz=None, xy=0.0, xz=0.0, yz=0.0, periodic=(True, True, True)):
        self.Lx = float(Lx)
        self.Ly = float(Lx if Ly is None else Ly)
        self.Lz = float(Lx if Lz is None else Lz)
        self.xy, self.xz, self.yz = float(xy), float(xz), float(yz)
        self.periodic = tuple(bool(p) for p in periodic)

    @classmethod
    def cube(cls, L):
        return cls(L, L, L)

    @property
    def L(self):
        return np.array([self.Lx, self.Ly, self.Lz])

    @property
    def volume(self):
        return self.Lx * self.Ly * self.Lz

    def as_abi(self):
        b = _abi.Box()
        b.Lx, b.Ly, b.Lz, b.xy, b.xz, b.yz = self.Lx, self.Ly, self.Lz, self.xy, self.xz, self.yz
        for d in range(3):
            b.periodic[d] = 1 if self.periodic[d] else 0
        return b


class Snapshot:
    """Host-side description of a configuration (the subset of ``hoomd.Snapshot`` the force path needs)."""

    def __init__(self, box, position, typeid=None, types=("A",), velocity=None, mass=None, orientation=None, tag=None,
                 bond_group=None, bond_typeid=None, bond_types=()):
        self.box = box
        self.position = np.ascontiguousarray(position, dtype=np.float64).reshape(-1, 3)
        n = self.N = self.position.shape[0]
        self.types = list(types)
        self.typeid = np.zeros(n, dtype=np.int64) if typeid is None else np.ascontiguousarray(typeid, dtype=np.int64)
        self.velocity = np.zeros((n, 3)) if velocity is None else np.ascontiguousarray(velocity, dtype=np.float64).reshape(-1, 3)
        self.mass = np.ones(n) if mass is None else np.ascontiguousarray(mass, dtype=np.float64)
        if orientation is None:
            orientation = np.tile(np.array([1.0, 0.0, 0.0, 0.0]), (n, 1))
        self.orientation = np.ascontiguousarray(orientation, dtype=np.float64).reshape(-1, 4)
        self.tag = np.arange(n, dtype=np.uint32) if tag is None else np.ascontiguousarray(tag, dtype=np.uint32)
        self.bond_group = np.zeros((0, 2), dtype=np.int64) if bond_group is None else np.ascontiguousarray(bond_group, dtype=np.int64).reshape(-1, 2)
        self.bond_typeid = (np.zeros(len(self.bond_group), dtype=np.int64) if bond_typeid is None
                            else np.ascontiguousarray(bond_typeid, dtype=np.int64))
        self.bond_types = list(bond_types) if len(bond_types) or len(self.bond_group) == 0 else ["A-A"]
        if len(self.bond_group) and not self.bond_types:
            self.bond_types = ["A-A"]
        for name, arr in (("typeid", self.typeid), ("mass", self.mass), ("tag", self.tag)):
            if arr.shape[0] != n:
                raise ValueError(f"{name} has {arr.shape[0]} entries for {n} particles")
        if self.typeid.size and (self.typeid.min() < 0 or self.typeid.max() >= len(self.types)):
            raise ValueError("typeid out of range")


def pack_scalar4(xyz, w_bits_or_vals, bits, w_is_int):
    """(n,3) + per-particle w -> (n,4) Scalar array; integer w is bit-cast like HOOMD's __int_as_scalar."""
    ftype, itype, _ = scalar_dtype(bits)
    n = xyz.shape[0]
    out = np.empty((n, 4), dtype=ftype)
    out[:, :3] = xyz
    if w_is_int:
        out[:, 3] = np.ascontiguousarray(w_bits_or_vals, dtype=itype).view(ftype)
    else:
        out[:, 3] = w_bits_or_vals
    return out


def build_bond_table(n_local, n_max, bond_group, bond_typeid):
    """Per-particle bond table in HOOMD's GPU layout.

    Returns (n_bonds[u32, n_local], bondlist[u32, max_bonds*pitch*2], bond_pos[u32, max_bonds*pitch], pitch, max_bonds);
    element (slot k, particle i) sits at ``k*pitch + i`` and holds (partner index, bond type). A bond is entered
    for each of its members that is local (index < n_local); the partner may be a ghost.
    """
    bond_group = np.asarray(bond_group, dtype=np.int64).reshape(-1, 2)
    bond_typeid = np.asarray(bond_typeid, dtype=np.int64)
    if len(bond_group) and (bond_group.min() < 0 or bond_group.max() >= n_max):
        raise ValueError("bond member index out of range (bond_group holds particle indices)")
    pitch = max(int(n_local), 1)
    members = np.concatenate([bond_group[:, 0], bond_group[:, 1]])
    partners = np.concatenate([bond_group[:, 1], bond_group[:, 0]])
    btypes = np.concatenate([bond_typeid, bond_typeid])
    posin = np.concatenate([np.zeros(len(bond_group), dtype=np.int64), np.ones(len(bond_group), dtype=np.int64)])
    keep = members < n_local
    members, partners, btypes, posin = members[keep], partners[keep], btypes[keep], posin[keep]
    order = np.argsort(members, kind="stable")
    members, partners, btypes, posin = members[order], partners[order], btypes[order], posin[order]
    n_bonds = np.bincount(members, minlength=n_local).astype(np.uint32)[:n_local] if n_local else np.zeros(0, np.uint32)
    max_bonds = int(n_bonds.max()) if n_bonds.size and n_bonds.max() > 0 else 1
    starts = np.zeros(n_local + 1, dtype=np.int64)
    np.cumsum(n_bonds, out=starts[1:])
    slot = np.arange(len(members)) - starts[members]
    bondlist = np.zeros((max_bonds, pitch, 2), dtype=np.uint32)
    bond_pos = np.zeros((max_bonds, pitch), dtype=np.uint32)
    bondlist[slot, members, 0] = partners
    bondlist[slot, members, 1] = btypes
    bond_pos[slot, members] = posin
    return n_bonds, bondlist.reshape(-1), bond_pos.reshape(-1), pitch, max_bonds


def build_bond_table_device(n_local, group, btype):
    """``build_bond_table`` on torch tensors (any device): ``group`` (nb, 2) int64 particle indices — a member may be a ghost
    (index >= n_local), ``btype`` (nb,) int64. Same layout and same slot order as the host version. Used where the table is
    rebuilt during a run (domain decomposition: bonds follow migrating particles, partners may be ghosts)."""
    dev = group.device
    nb = group.shape[0]
    pitch = max(int(n_local), 1)
    members = torch.cat([group[:, 0], group[:, 1]])
    partners = torch.cat([group[:, 1], group[:, 0]])
    types = torch.cat([btype, btype])
    posin = torch.cat([torch.zeros(nb, dtype=torch.int64, device=dev), torch.ones(nb, dtype=torch.int64, device=dev)])
    keep = members < n_local
    members, partners, types, posin = members[keep], partners[keep], types[keep], posin[keep]
    order = torch.argsort(members, stable=True)
    members, partners, types, posin = members[order], partners[order], types[order], posin[order]
    n_bonds = torch.bincount(members, minlength=pitch)[:pitch] if members.numel() else torch.zeros(pitch, dtype=torch.int64, device=dev)
    max_bonds = max(int(n_bonds.max().item()), 1) if n_bonds.numel() else 1
    starts = torch.cumsum(n_bonds, 0) - n_bonds
    slot = torch.arange(members.numel(), device=dev) - starts[members]
    bondlist = torch.zeros((max_bonds, pitch, 2), dtype=torch.int32, device=dev)
    bond_pos = torch.zeros((max_bonds, pitch), dtype=torch.int32, device=dev)
    bondlist[slot, members, 0] = partners.to(torch.int32)
    bondlist[slot, members, 1] = types.to(torch.int32)
    bond_pos[slot, members] = posin.to(torch.int32)
    n_out = n_bonds[:n_local].to(torch.int32) if n_local else torch.zeros(1, dtype=torch.int32, device=dev)
    return n_out, bondlist.reshape(-1), bond_pos.reshape(-1), pitch, max_bonds


class State:
    """Device-resident particle data of one rank: ``N`` local particles followed by ``n_ghost`` ghosts."""

    def __init__(self, snapshot, device="cuda", scalar_bits=32, seed=0):
        self.scalar_bits = int(scalar_bits)
        _, _, self.tdtype = scalar_dtype(self.scalar_bits)
        self.device = torch.device(device)
        self.box = snapshot.box
        self.types = list(snapshot.types)
        self.bond_types = list(snapshot.bond_types)
        self.N = snapshot.N
        self.n_ghost = 0
        self.seed = int(seed) & 0xFFFF  # HOOMD simulation seed is 16 bit
        s = snapshot
        self.pos = self._to_dev(pack_scalar4(s.position, s.typeid, self.scalar_bits, True))
        self.vel = self._to_dev(pack_scalar4(s.velocity, s.mass, self.scalar_bits, False))
        self.orientation = self._to_dev(s.orientation.astype(scalar_dtype(self.scalar_bits)[0]))
        self.tag = self._to_dev(s.tag.astype(np.uint32).view(np.int32))
        self.typeid_host = s.typeid.copy()
        self.bond_group = s.bond_group.copy()  # particle indices
        self.bond_typeid = s.bond_typeid.copy()
        # the same bonds by particle TAG: the identity that survives migration between domains (azplugins_b200/domain.py)
        self.bond_tags = s.tag[s.bond_group].astype(np.uint32) if len(s.bond_group) else np.zeros((0, 2), dtype=np.uint32)
        self._bond_table = None
        self._bonds_dev = None  # (group int64 (nb, 2), type int64 (nb,)) device tensors of LOCAL indices, set by the domain layer
        self.image = None

    @property
    def n_max(self):
        return self.N + self.n_ghost

    def _to_dev(self, arr):
        return torch.from_numpy(np.ascontiguousarray(arr)).to(self.device)

    def set_local_bonds(self, group, btype):
        """Bonds of this rank as device tensors of local particle indices (a partner may be a ghost). Called by the domain layer
        after every migration / ghost re-selection; invalidates the per-particle table and the exclusion list."""
        self._bonds_dev = (group, btype)
        self._bond_table = None

    @property
    def has_bonds(self):
        return bool(self._bonds_dev[0].shape[0]) if self._bonds_dev is not None else bool(len(self.bond_group))

    def bond_table(self):
        """Device copy of the per-particle bond table (rebuilt when the domain layer hands in new local bonds)."""
        if self._bond_table is None:
            if self._bonds_dev is not None:
                n_bonds, bondlist, bond_pos, pitch, max_bonds = build_bond_table_device(self.N, *self._bonds_dev)
                self._bond_table = dict(n_bonds=n_bonds, bondlist=bondlist, bond_pos=bond_pos, pitch=pitch, max_bonds=max_bonds)
            else:
                n_bonds, bondlist, bond_pos, pitch, max_bonds = build_bond_table(self.N, self.n_max, self.bond_group, self.bond_typeid)
                self._bond_table = dict(
                    n_bonds=self._to_dev(n_bonds.view(np.int32)), bondlist=self._to_dev(bondlist.view(np.int32)),
                    bond_pos=self._to_dev(bond_pos.view(np.int32)), pitch=pitch, max_bonds=max_bonds)
        return self._bond_table

    def exclusions_device(self):
        """Bond exclusions as device CSR (head int64 [N + 1], list int32) over local particles."""
        if self._bonds_dev is None:
            head, lst = self.exclusions_csr()
            return (torch.from_numpy(head.view(np.int64)).to(self.device),
                    torch.from_numpy(lst.view(np.int32) if lst.size else np.zeros(1, np.int32)).to(self.device))
        g = self._bonds_dev[0]
        a = torch.cat([g[:, 0], g[:, 1]])
        b = torch.cat([g[:, 1], g[:, 0]])
        keep = a < self.N
        a, b = a[keep], b[keep]
        order = torch.argsort(a, stable=True)
        a, b = a[order], b[order]
        head = torch.zeros(self.N + 1, dtype=torch.int64, device=g.device)
        if a.numel():
            head[1:] = torch.cumsum(torch.bincount(a, minlength=self.N)[: self.N], 0)
        lst = b.to(torch.int32) if b.numel() else torch.zeros(1, dtype=torch.int32, device=g.device)
        return head, lst.contiguous()

    def exclusions_csr(self):
        """Bond exclusions (HOOMD nlist default ``exclusions=('bond',)``) as CSR over local particles."""
        g = self.bond_group
        a = np.concatenate([g[:, 0], g[:, 1]])
        b = np.concatenate([g[:, 1], g[:, 0]])
        keep = a < self.N
        a, b = a[keep], b[keep]
        order = np.argsort(a, kind="stable")
        a, b = a[order], b[order]
        head = np.zeros(self.N + 1, dtype=np.uint64)
        np.cumsum(np.bincount(a, minlength=self.N)[: self.N], out=head[1:])
        return head, b.astype(np.uint32)
