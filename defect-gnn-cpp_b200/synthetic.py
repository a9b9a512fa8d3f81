"""Seeded synthetic workloads for the BASELINE.json configs.

Everything is fp64 numpy; structures are returned as a `Batch` in the SoA layout the C ABI takes
(include/dgnn_cabi.h): atom_offsets[int64 S+1], x/y/z[float64 N_tot] Cartesian, lattices[float64 S*9] rows a,b,c.
Cartesian positions follow crystal::Structure (reference src/crystal/structure.cpp:17): pos = L^T f.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np


@dataclass
class Batch:
    atom_offsets: np.ndarray  # int64 [S+1]
    x: np.ndarray
    y: np.ndarray
    z: np.ndarray
    lattices: np.ndarray      # float64 [S, 9]
    species: np.ndarray | None = None  # int32 [N_tot] POSCAR species index (Atom::element)
    frac: np.ndarray | None = None     # float64 [N_tot, 3] fractional coordinates the positions were made from (the
                                       # reference's crystal::Structure is built from these, structure.cpp:13-19)

    @property
    def n_structs(self) -> int:
        return int(self.atom_offsets.shape[0] - 1)

    @property
    def n_atoms(self) -> int:
        return int(self.atom_offsets[-1])

    def structure(self, s: int):
        a0, a1 = int(self.atom_offsets[s]), int(self.atom_offsets[s + 1])
        return self.x[a0:a1], self.y[a0:a1], self.z[a0:a1], self.lattices[s].reshape(3, 3)

    def slice(self, s0: int, s1: int) -> "Batch":
        a0, a1 = int(self.atom_offsets[s0]), int(self.atom_offsets[s1])
        return Batch(self.atom_offsets[s0:s1 + 1] - a0, self.x[a0:a1], self.y[a0:a1], self.z[a0:a1],
                     self.lattices[s0:s1], None if self.species is None else self.species[a0:a1],
                     None if self.frac is None else self.frac[a0:a1])


def frac_to_cart(frac: np.ndarray, lattice: np.ndarray):
    """pos_k = (L[0,k] f0 + L[1,k] f1) + L[2,k] f2 -- the order Eigen 3.4 evaluates lattice.transpose() * frac in on
    x86-64 (vectorised redux over a contiguous column); pinned on the reference's structure.cpp:17 compiled verbatim
    (oracle/_ref, tests/test_ref_pin.py)."""
    L = np.asarray(lattice, dtype=np.float64).reshape(3, 3)
    f = np.asarray(frac, dtype=np.float64).reshape(-1, 3)
    out = []
    for k in range(3):
        out.append((L[0, k] * f[:, 0] + L[1, k] * f[:, 1]) + L[2, k] * f[:, 2])
    return out[0], out[1], out[2]


def concat(structs) -> Batch:
    """structs: iterable of (x, y, z, lattice[3x3]), (x, y, z, lattice, species) or (..., species, frac)."""
    xs, ys, zs, lats, sp, fr, offs = [], [], [], [], [], [], [0]
    have_sp = True
    have_fr = True
    for st in structs:
        x, y, z, lat = st[:4]
        xs.append(np.asarray(x, dtype=np.float64))
        ys.append(np.asarray(y, dtype=np.float64))
        zs.append(np.asarray(z, dtype=np.float64))
        lats.append(np.asarray(lat, dtype=np.float64).reshape(9))
        offs.append(offs[-1] + xs[-1].shape[0])
        if len(st) > 4 and st[4] is not None:
            sp.append(np.asarray(st[4], dtype=np.int32))
        else:
            have_sp = False
        if len(st) > 5 and st[5] is not None:
            fr.append(np.asarray(st[5], dtype=np.float64).reshape(-1, 3))
        else:
            have_fr = False
    cat = (lambda v: np.concatenate(v) if v else np.zeros(0, dtype=np.float64))
    return Batch(np.asarray(offs, dtype=np.int64), cat(xs), cat(ys), cat(zs),
                 np.stack(lats) if lats else np.zeros((0, 9), dtype=np.float64),
                 np.concatenate(sp) if (have_sp and sp) else None,
                 np.concatenate(fr) if (have_fr and fr) else None)


# ---- C2 / C4 / C5: cubic ABO3 perovskite supercells with one vacancy ------------------------------------------
_BASIS = np.array([[0.0, 0.0, 0.0], [0.5, 0.5, 0.5], [0.5, 0.5, 0.0], [0.5, 0.0, 0.5], [0.0, 0.5, 0.5]])
_BASIS_SPECIES = np.array([0, 1, 2, 2, 2], dtype=np.int32)


def perovskite_supercell(nx: int, ny: int, nz: int, seed: int, a: float = 4.0, jitter: float = 0.02,
                         vacancy: bool = True, with_frac: bool = False):
    """One nx x ny x nz ABO3 supercell: Gaussian jitter (sigma Angstrom per coordinate), one random site deleted.

    Returns (x, y, z, lattice[3x3], species) -- plus frac with with_frac=True.  Fractional coordinates are NOT wrapped (jitter can push sites at
    frac 0 slightly negative), like real relaxed POSCARs and like the reference, which never wraps.
    """
    rng = np.random.Generator(np.random.PCG64(seed))
    ix, iy, iz = np.meshgrid(np.arange(nx), np.arange(ny), np.arange(nz), indexing="ij")
    cells = np.stack([ix.ravel(), iy.ravel(), iz.ravel()], axis=1).astype(np.float64)  # [C,3]
    ideal = (cells[:, None, :] + _BASIS[None, :, :]).reshape(-1, 3) * a                   # [5C,3] Angstrom
    species = np.tile(_BASIS_SPECIES, cells.shape[0])
    pos = ideal + rng.normal(0.0, jitter, size=ideal.shape)
    if vacancy:
        kill = int(rng.integers(0, pos.shape[0]))
        keep = np.ones(pos.shape[0], dtype=bool)
        keep[kill] = False
        pos, species = pos[keep], species[keep]
    lengths = np.array([nx * a, ny * a, nz * a])
    lattice = np.diag(lengths)
    frac = pos / lengths[None, :]
    x, y, z = frac_to_cart(frac, lattice)
    return (x, y, z, lattice, species, frac) if with_frac else (x, y, z, lattice, species)


def c2_batch(n_structs: int = 10_000, seed0: int = 1234, dims=(4, 4, 2)) -> Batch:
    """BASELINE config 2: n_structs x (4x4x2 ABO3, a=4.0 -> 16x16x8 Angstrom, 159 atoms); seed = seed0 + s."""
    return concat(perovskite_supercell(*dims, seed=seed0 + s, with_frac=True) for s in range(n_structs))


def c4_structure(dims=(59, 59, 58), seed: int = 4321) -> Batch:
    """BASELINE config 4: one ~1M-atom perovskite supercell (59x59x58 cells -> 1,009,489 atoms)."""
    return concat([perovskite_supercell(*dims, seed=seed, with_frac=True)])


# ---- C3: small triclinic cells with short lattice vectors -----------------------------------------------------
def lattice_from_params(a, b, c, alpha_deg, beta_deg, gamma_deg) -> np.ndarray:
    al, be, ga = np.deg2rad([alpha_deg, beta_deg, gamma_deg])
    va = np.array([a, 0.0, 0.0])
    vb = np.array([b * np.cos(ga), b * np.sin(ga), 0.0])
    cx = c * np.cos(be)
    cy = c * (np.cos(al) - np.cos(be) * np.cos(ga)) / np.sin(ga)
    cz2 = c * c - cx * cx - cy * cy
    if cz2 <= 0:
        raise ValueError("invalid cell angles")
    return np.stack([va, vb, np.array([cx, cy, np.sqrt(cz2)])])


def triclinic_cell(seed: int, len_range=(3.0, 6.0), ang_range=(60.0, 120.0), n_atoms_range=(2, 8),
                   min_dist: float = 1.0, with_frac: bool = False):
    rng = np.random.Generator(np.random.PCG64(seed))
    while True:
        a, b, c = rng.uniform(*len_range, size=3)
        al, be, ga = rng.uniform(*ang_range, size=3)
        try:
            L = lattice_from_params(a, b, c, al, be, ga)
        except ValueError:
            continue
        if abs(np.linalg.det(L)) >= 0.5 * a * b * c:
            break
    n = int(rng.integers(n_atoms_range[0], n_atoms_range[1] + 1))
    shifts = np.array([[i, j, k] for i in (-1, 0, 1) for j in (-1, 0, 1) for k in (-1, 0, 1)], dtype=np.float64)
    fr = []
    tries = 0
    while len(fr) < n and tries < 10_000:
        tries += 1
        f = rng.uniform(0.0, 1.0, size=3)
        ok = True
        for g in fr:
            d = ((f - g)[None, :] + shifts) @ L
            if np.min(np.einsum("ij,ij->i", d, d)) < min_dist * min_dist:
                ok = False
                break
        if ok:
            fr.append(f)
    frac = np.array(fr)
    x, y, z = frac_to_cart(frac, L)
    sp = np.zeros(frac.shape[0], dtype=np.int32)
    return (x, y, z, L, sp, frac) if with_frac else (x, y, z, L, sp)


def c3_batch(n_structs: int = 10_000, seed0: int = 777) -> Batch:
    return concat(triclinic_cell(seed0 + s, with_frac=True) for s in range(n_structs))


def adversarial_cells() -> Batch:
    """Strongly skewed cells (gamma = 15 / 30 degrees) where the reference's image cube (from vector norms,
    neighbor_list.cpp:68-72) misses true neighbours; the product must reproduce the reference range, not the
    exact one (rule R3).  Plus tiny / needle cells with many self-images."""
    out = []
    rng = np.random.Generator(np.random.PCG64(99))
    for gamma in (15.0, 30.0):
        L = lattice_from_params(3.0, 4.0, 5.0, 70.0, 80.0, gamma) if gamma > 20 else \
            lattice_from_params(3.0, 4.0, 5.0, 85.0, 80.0, gamma)
        frac = rng.uniform(0, 1, size=(4, 3))
        x, y, z = frac_to_cart(frac, L)
        out.append((x, y, z, L, np.zeros(4, dtype=np.int32), frac))
    # single atom in a 2.5 Angstrom cube: pure self-image edges
    L = np.diag([2.5, 2.5, 2.5])
    frac = np.array([[0.1, 0.2, 0.3]])
    x, y, z = frac_to_cart(frac, L)
    out.append((x, y, z, L, np.zeros(1, dtype=np.int32), frac))
    # needle cell 2 x 2 x 30 with atoms given far outside [0,1) in fractional coordinates (never wrapped)
    L = np.diag([2.0, 2.0, 30.0])
    frac = np.array([[0.0, 0.0, 0.0], [1.25, -0.5, 0.5], [0.5, 2.5, -0.25]])
    x, y, z = frac_to_cart(frac, L)
    out.append((x, y, z, L, np.zeros(3, dtype=np.int32), frac))
    # two atoms at the identical position (kept with d = 0 because idx differs), in a rotated cubic cell
    th = np.deg2rad(33.0)
    Rz = np.array([[np.cos(th), -np.sin(th), 0], [np.sin(th), np.cos(th), 0], [0, 0, 1.0]])
    L = (np.diag([5.0, 5.0, 5.0]) @ Rz.T)
    frac = np.array([[0.25, 0.25, 0.25], [0.25, 0.25, 0.25], [0.75, 0.5, 0.1]])
    x, y, z = frac_to_cart(frac, L)
    out.append((x, y, z, L, np.zeros(3, dtype=np.int32), frac))
    return concat(out)


def simple_cubic(a: float = 1.0, n: int = 1) -> Batch:
    """n x n x n simple-cubic supercell, lattice constant a (known-answer shells 6,12,8,6,24,24,12,30)."""
    ix, iy, iz = np.meshgrid(np.arange(n), np.arange(n), np.arange(n), indexing="ij")
    frac = np.stack([ix.ravel(), iy.ravel(), iz.ravel()], axis=1).astype(np.float64) / n
    L = np.diag([a * n] * 3)
    x, y, z = frac_to_cart(frac, L)
    return concat([(x, y, z, L, np.zeros(frac.shape[0], dtype=np.int32), frac)])


# ---- POSCAR text -> arrays (used only to load fixtures; the reference io layer is out of scope) ----------------
def parse_poscar(text: str):
    """Same field semantics as reference src/io/vasp_parser.cpp:13-78 (scale, rows a/b/c, Direct|Cartesian)."""
    lines = text.splitlines()
    scale = float(lines[1].split()[0])
    L = np.array([[float(t) for t in lines[2 + i].split()[:3]] for i in range(3)]) * scale
    counts = [int(t) for t in lines[6].split()]
    direct = lines[7].strip()[:1] in ("d", "D")
    n = sum(counts)
    coords = np.array([[float(t) for t in lines[8 + i].split()[:3]] for i in range(n)])
    species = np.concatenate([np.full(c, k, dtype=np.int32) for k, c in enumerate(counts)])
    frac = coords if direct else coords @ np.linalg.inv(L)
    return L, frac, species
