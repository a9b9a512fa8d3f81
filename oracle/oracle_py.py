"""ctypes binding to the CPU oracle (oracle/libdgnn_oracle.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs.  The product package never imports this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libdgnn_oracle.so")

ALGO_BRUTE, ALGO_KDTREE, ALGO_CELLLIST = 0, 1, 2
UNCAPPED = 2**64 - 1  # std::numeric_limits<size_t>::max() (betti_features.cpp:107)


class _OrcGraph(C.Structure):
    _fields_ = [
        ("n_atoms", C.c_int64),
        ("n_edges", C.c_int64),
        ("row_ptr", C.POINTER(C.c_int64)),
        ("dst", C.POINTER(C.c_int32)),
        ("image", C.POINTER(C.c_int8)),
        ("dist", C.POINTER(C.c_double)),
        ("disp", C.POINTER(C.c_double)),
        ("n_near_cutoff", C.c_int64),
        ("n_tie_rows", C.c_int64),
        ("num_images", C.c_int32),
    ]


class _BatchResult(C.Structure):
    _fields_ = [("n_edges", C.c_int64), ("checksum", C.c_double), ("seconds", C.c_double), ("threads", C.c_int32)]


def build(force: bool = False) -> str:
    """Compile the oracle with its Makefile (gcc only)."""
    src = os.path.join(HERE, "dgnn_oracle.cpp")
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", HERE, "all"], stdout=subprocess.DEVNULL)
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(LIB_PATH)
        dp = C.POINTER(C.c_double)
        L.orc_num_images.argtypes = [dp, C.c_double]
        L.orc_num_images.restype = C.c_int
        L.orc_neighbor_list.argtypes = [C.c_int64, dp, dp, dp, dp, C.c_double, C.c_uint64, C.c_double, C.c_int,
                                        C.POINTER(_OrcGraph)]
        L.orc_neighbor_list.restype = C.c_int
        L.orc_free_graph.argtypes = [C.POINTER(_OrcGraph)]
        L.orc_gaussian_rbf.argtypes = [C.c_double, C.c_double, C.c_double, dp]
        L.orc_gaussian_rbf.restype = C.c_int
        L.orc_edge_attr_colmajor.argtypes = [C.c_int64, dp, C.c_double, C.c_double, dp]
        L.orc_edge_attr_colmajor.restype = C.c_int
        L.orc_edge_attr_rowmajor.argtypes = [C.c_int64, dp, C.c_double, C.c_double, dp]
        L.orc_edge_attr_rowmajor.restype = C.c_int
        L.orc_frac_to_cart.argtypes = [C.c_int64, dp, dp, dp, dp, dp]
        i64p = C.POINTER(C.c_int64)
        L.orc_build_batch.argtypes = [C.c_int64, i64p, dp, dp, dp, dp, C.c_double, C.c_uint64, C.c_double,
                                      C.c_double, C.c_double, C.c_int, C.c_int, C.POINTER(_BatchResult)]
        L.orc_build_batch.restype = C.c_int
        L.orc_batch_edge_counts.argtypes = [C.c_int64, i64p, dp, dp, dp, dp, C.c_double, C.c_uint64, C.c_double,
                                            C.c_int, C.c_int, i64p]
        L.orc_batch_edge_counts.restype = C.c_int
        L.orc_rows_brute.argtypes = [C.c_int64, dp, dp, dp, dp, C.c_double, C.c_double, C.c_int64, i64p,
                                     C.c_int64, i64p, C.POINTER(C.c_int32), C.POINTER(C.c_int8), dp]
        L.orc_rows_brute.restype = C.c_int
        _lib = L
    return _lib


def _dp(a: np.ndarray):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _f64(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.float64)


@dataclass
class OracleGraph:
    row_ptr: np.ndarray  # int64 [N+1]
    dst: np.ndarray      # int32 [E]
    image: np.ndarray    # int8  [E,4]  (n_a, n_b, n_c, 0)
    dist: np.ndarray     # float64 [E]
    disp: np.ndarray     # float64 [E,3]
    n_near_cutoff: int
    n_tie_rows: int
    num_images: int

    @property
    def n_edges(self) -> int:
        return int(self.dst.shape[0])


def num_images(lattice, r_cutoff: float) -> int:
    Lm = _f64(lattice).reshape(9)
    return int(lib().orc_num_images(_dp(Lm), float(r_cutoff)))


def neighbor_list(x, y, z, lattice, r_cutoff=10.0, max_neighbors=20, epsilon=1e-10, algo=ALGO_BRUTE) -> OracleGraph:
    x, y, z = _f64(x), _f64(y), _f64(z)
    Lm = _f64(lattice).reshape(9)
    g = _OrcGraph()
    rc = lib().orc_neighbor_list(x.shape[0], _dp(x), _dp(y), _dp(z), _dp(Lm), float(r_cutoff),
                                 C.c_uint64(int(max_neighbors)), float(epsilon), int(algo), C.byref(g))
    if rc != 0:
        raise RuntimeError(f"orc_neighbor_list failed rc={rc}")
    n, e = g.n_atoms, g.n_edges
    out = OracleGraph(
        row_ptr=np.ctypeslib.as_array(g.row_ptr, shape=(n + 1,)).copy(),
        dst=np.ctypeslib.as_array(g.dst, shape=(max(e, 1),))[:e].copy(),
        image=np.ctypeslib.as_array(g.image, shape=(max(e, 1) * 4,))[: e * 4].copy().reshape(e, 4),
        dist=np.ctypeslib.as_array(g.dist, shape=(max(e, 1),))[:e].copy(),
        disp=np.ctypeslib.as_array(g.disp, shape=(max(e, 1) * 3,))[: e * 3].copy().reshape(e, 3),
        n_near_cutoff=int(g.n_near_cutoff), n_tie_rows=int(g.n_tie_rows), num_images=int(g.num_images))
    lib().orc_free_graph(C.byref(g))
    return out


def rbf_bins(r_cutoff: float, dr: float) -> int:
    return int(lib().orc_gaussian_rbf(0.0, float(r_cutoff), float(dr), None))


def gaussian_rbf(distance: float, r_cutoff: float, dr: float) -> np.ndarray:
    k = rbf_bins(r_cutoff, dr)
    out = np.empty(k, dtype=np.float64)
    lib().orc_gaussian_rbf(float(distance), float(r_cutoff), float(dr), _dp(out))
    return out


def edge_attr(dist, r_cutoff: float, dr: float, colmajor: bool = False) -> np.ndarray:
    """E x K fp64 features. colmajor=True returns the reference's storage order as a Fortran-ordered array."""
    dist = _f64(dist)
    k = rbf_bins(r_cutoff, dr)
    e = dist.shape[0]
    if colmajor:
        out = np.empty(e * k, dtype=np.float64)
        lib().orc_edge_attr_colmajor(e, _dp(dist), float(r_cutoff), float(dr), _dp(out))
        return out.reshape(k, e).T
    out = np.empty((e, k), dtype=np.float64)
    lib().orc_edge_attr_rowmajor(e, _dp(dist), float(r_cutoff), float(dr), _dp(out))
    return out


def frac_to_cart(frac, lattice):
    frac = _f64(frac).reshape(-1, 3)
    Lm = _f64(lattice).reshape(9)
    n = frac.shape[0]
    x, y, z = (np.empty(n, dtype=np.float64) for _ in range(3))
    lib().orc_frac_to_cart(n, _dp(frac), _dp(Lm), _dp(x), _dp(y), _dp(z))
    return x, y, z


def build_batch(atom_offsets, x, y, z, lattices, r_cutoff, max_neighbors, epsilon, rbf_r_cutoff, dr,
                algo=ALGO_KDTREE, threads=1):
    """Reference-style CPU build of a whole batch; returns (n_edges, checksum, seconds, threads)."""
    ao = np.ascontiguousarray(atom_offsets, dtype=np.int64)
    x, y, z, lat = _f64(x), _f64(y), _f64(z), _f64(lattices).reshape(-1)
    res = _BatchResult()
    rc = lib().orc_build_batch(ao.shape[0] - 1, ao.ctypes.data_as(C.POINTER(C.c_int64)), _dp(x), _dp(y), _dp(z),
                               _dp(lat), float(r_cutoff), C.c_uint64(int(max_neighbors)), float(epsilon),
                               float(rbf_r_cutoff), float(dr), int(algo), int(threads), C.byref(res))
    if rc != 0:
        raise RuntimeError(f"orc_build_batch failed rc={rc}")
    return int(res.n_edges), float(res.checksum), float(res.seconds), int(res.threads)


def batch_edge_counts(atom_offsets, x, y, z, lattices, r_cutoff, max_neighbors=UNCAPPED, epsilon=1e-10,
                      algo=ALGO_CELLLIST, threads=0) -> np.ndarray:
    ao = np.ascontiguousarray(atom_offsets, dtype=np.int64)
    x, y, z, lat = _f64(x), _f64(y), _f64(z), _f64(lattices).reshape(-1)
    counts = np.zeros(ao.shape[0] - 1, dtype=np.int64)
    i64p = C.POINTER(C.c_int64)
    lib().orc_batch_edge_counts(ao.shape[0] - 1, ao.ctypes.data_as(i64p), _dp(x), _dp(y), _dp(z), _dp(lat),
                                float(r_cutoff), C.c_uint64(int(max_neighbors)), float(epsilon), int(algo),
                                int(threads), counts.ctypes.data_as(i64p))
    return counts


def rows_brute(x, y, z, lattice, r_cutoff, sources, row_cap, epsilon=1e-10):
    x, y, z = _f64(x), _f64(y), _f64(z)
    Lm = _f64(lattice).reshape(9)
    src = np.ascontiguousarray(sources, dtype=np.int64)
    q = src.shape[0]
    deg = np.zeros(q, dtype=np.int64)
    dst = np.full((q, row_cap), -1, dtype=np.int32)
    image = np.zeros((q, row_cap, 4), dtype=np.int8)
    dist = np.zeros((q, row_cap), dtype=np.float64)
    i64p = C.POINTER(C.c_int64)
    rc = lib().orc_rows_brute(x.shape[0], _dp(x), _dp(y), _dp(z), _dp(Lm), float(r_cutoff), float(epsilon), q,
                              src.ctypes.data_as(i64p), int(row_cap), deg.ctypes.data_as(i64p),
                              dst.ctypes.data_as(C.POINTER(C.c_int32)), image.ctypes.data_as(C.POINTER(C.c_int8)),
                              _dp(dist))
    if rc != 0:
        raise RuntimeError(f"orc_rows_brute failed rc={rc}")
    return deg, dst, image, dist
