"""ctypes binding to oracle/_ref/libref_graph.so -- the reference's OWN graph-layer sources compiled verbatim
(oracle/Makefile target `ref`, driver oracle/ref_driver.cpp, third-party headers replaced by oracle/ref_shim/).

TEST INFRASTRUCTURE ONLY: imported by tests/ (to pin the oracle restatement and the golden fixtures on reference code)
and by bench.py's cpu_baseline / --impl reference legs.  The product package never imports this module.
The library is built in the container (where /root/reference exists) and travels to the GPU box prebuilt.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "_ref", "libref_graph.so")
REF_ROOT = "/root/reference"
UNCAPPED = 2**64 - 1


def build(force: bool = False) -> str | None:
    """(Re)build from /root/reference when it is present; otherwise use the prebuilt library if there is one."""
    if os.path.isdir(os.path.join(REF_ROOT, "src", "graph")):
        deps = [os.path.join(HERE, "ref_driver.cpp"), os.path.join(HERE, "ref_shim", "Eigen", "Dense"),
                os.path.join(HERE, "ref_shim", "nanoflann.hpp"), os.path.join(HERE, "Makefile")]
        if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < max(os.path.getmtime(d) for d in deps):
            subprocess.check_call(["make", "-C", HERE, "ref"], stdout=subprocess.DEVNULL)
    return LIB_PATH if os.path.exists(LIB_PATH) else None


def available() -> bool:
    return build() is not None


_lib = None


class _BatchResult(C.Structure):
    _fields_ = [("n_edges", C.c_int64), ("checksum", C.c_double), ("seconds", C.c_double), ("threads", C.c_int32)]


def lib():
    global _lib
    if _lib is None:
        path = build()
        if path is None:
            raise RuntimeError("oracle/_ref/libref_graph.so is missing and /root/reference is not present to build it")
        L = C.CDLL(path)
        dp, i64p, i32p = C.POINTER(C.c_double), C.POINTER(C.c_int64), C.POINTER(C.c_int32)
        L.ref_last_error.restype = C.c_char_p
        L.ref_nl_create.argtypes = [C.c_int64, dp, dp, i32p, C.c_double, C.c_uint64, C.c_double]
        L.ref_nl_create.restype = C.c_void_p
        L.ref_nl_num_atoms.argtypes = [C.c_void_p]
        L.ref_nl_num_atoms.restype = C.c_int64
        L.ref_nl_num_edges.argtypes = [C.c_void_p]
        L.ref_nl_num_edges.restype = C.c_int64
        L.ref_nl_export.argtypes = [C.c_void_p, dp, i64p, i64p, dp, dp]
        L.ref_nl_free.argtypes = [C.c_void_p]
        L.ref_gaussian_rbf.argtypes = [C.c_double, C.c_double, C.c_double, dp]
        L.ref_gaussian_rbf.restype = C.c_int
        L.ref_graph_create.argtypes = [C.c_void_p, C.c_int, i32p, dp, C.c_int, C.c_double, C.c_double]
        L.ref_graph_create.restype = C.c_void_p
        L.ref_graph_dims.argtypes = [C.c_void_p, i64p, i64p, i64p, i64p]
        L.ref_graph_export.argtypes = [C.c_void_p, dp, i32p, dp, dp]
        L.ref_graph_free.argtypes = [C.c_void_p]
        L.ref_parse_vasp.argtypes = [C.c_char_p, dp, dp, i32p, C.c_int64]
        L.ref_parse_vasp.restype = C.c_int64
        L.ref_build_batch.argtypes = [C.c_int64, i64p, dp, dp, i32p, C.c_double, C.c_uint64, C.c_double, C.c_double,
                                      C.c_double, C.c_int, C.c_int, C.POINTER(_BatchResult)]
        L.ref_build_batch.restype = C.c_int
        _lib = L
    return _lib


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


@dataclass
class RefNeighborList:
    """What graph::NeighborList holds, read out of the reference's object (rows in the reference's own order)."""
    pos: np.ndarray       # float64 [N,3]  Structure::atoms()[i].position
    row_ptr: np.ndarray   # int64 [N+1]
    idx: np.ndarray       # int64 [E]      Neighbor::idx
    dist: np.ndarray      # float64 [E]    Neighbor::distance
    disp: np.ndarray      # float64 [E,3]  Neighbor::displacement

    @property
    def n_edges(self):
        return int(self.idx.shape[0])


@dataclass
class RefGraph:
    node_features: np.ndarray  # float64 [N,F]   (read from column-major storage)
    edge_index: np.ndarray     # int32 [2,E]
    edge_attr: np.ndarray      # float64 [E,K]
    edge_index_raw: np.ndarray  # the reference's memory order [src0,dst0,src1,dst1,...]
    edge_attr_raw: np.ndarray   # bin-major E x K doubles
    target: float


class Structure:
    """Handle on a reference crystal::Structure + graph::NeighborList pair."""

    def __init__(self, frac, lattice, atom_types=None, r_cutoff=10.0, max_neighbors=20, epsilon=1e-10):
        frac = _f64(frac).reshape(-1, 3)
        lat = _f64(lattice).reshape(9)
        at = None if atom_types is None else np.ascontiguousarray(atom_types, dtype=np.int32)
        self._h = lib().ref_nl_create(frac.shape[0], _dp(frac), _dp(lat),
                                      None if at is None else at.ctypes.data_as(C.POINTER(C.c_int32)),
                                      float(r_cutoff), C.c_uint64(int(max_neighbors)), float(epsilon))
        if not self._h:
            raise RuntimeError("reference NeighborList failed: " + lib().ref_last_error().decode())

    def neighbor_list(self) -> RefNeighborList:
        L = lib()
        n, e = L.ref_nl_num_atoms(self._h), L.ref_nl_num_edges(self._h)
        pos = np.empty((n, 3)); rp = np.empty(n + 1, np.int64); idx = np.empty(max(e, 1), np.int64)
        dist = np.empty(max(e, 1)); disp = np.empty((max(e, 1), 3))
        L.ref_nl_export(self._h, _dp(pos), rp.ctypes.data_as(C.POINTER(C.c_int64)),
                        idx.ctypes.data_as(C.POINTER(C.c_int64)), _dp(dist), _dp(disp))
        return RefNeighborList(pos, rp, idx[:e], dist[:e], disp[:e])

    def crystal_graph(self, embeddings: dict, dims=92, r_cutoff=10.0, dr=0.1) -> RefGraph:
        """graph::CrystalGraph; raises KeyError where the reference's std::map::at throws std::out_of_range."""
        L = lib()
        keys = np.ascontiguousarray(sorted(embeddings), dtype=np.int32)
        table = _f64(np.stack([np.asarray(embeddings[int(k)], dtype=np.float64) for k in keys])) if len(keys) else \
            np.zeros((0, dims))
        g = L.ref_graph_create(self._h, len(keys), keys.ctypes.data_as(C.POINTER(C.c_int32)), _dp(table), int(dims),
                               float(r_cutoff), float(dr))
        if not g:
            raise KeyError(L.ref_last_error().decode())
        try:
            nn, ne, f, k = (C.c_int64() for _ in range(4))
            L.ref_graph_dims(g, C.byref(nn), C.byref(ne), C.byref(f), C.byref(k))
            nn, ne, f, k = nn.value, ne.value, f.value, k.value
            nf = np.empty(max(nn * f, 1)); ei = np.empty(max(2 * ne, 1), np.int32); ea = np.empty(max(ne * k, 1))
            tgt = C.c_double()
            L.ref_graph_export(g, _dp(nf), ei.ctypes.data_as(C.POINTER(C.c_int32)), _dp(ea), C.byref(tgt))
            nf, ei, ea = nf[:nn * f], ei[:2 * ne], ea[:ne * k]
            return RefGraph(nf.reshape(f, nn).T.copy(), ei.reshape(ne, 2).T.copy(), ea.reshape(k, ne).T.copy(),
                            ei.copy(), ea.copy(), tgt.value)
        finally:
            L.ref_graph_free(g)

    def close(self):
        if self._h:
            lib().ref_nl_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def neighbor_list(frac, lattice, r_cutoff=10.0, max_neighbors=20, epsilon=1e-10) -> RefNeighborList:
    s = Structure(frac, lattice, None, r_cutoff, max_neighbors, epsilon)
    try:
        return s.neighbor_list()
    finally:
        s.close()


def gaussian_rbf(distance, r_cutoff, dr) -> np.ndarray:
    k = lib().ref_gaussian_rbf(0.0, float(r_cutoff), float(dr), None)
    out = np.empty(max(k, 1))
    lib().ref_gaussian_rbf(float(distance), float(r_cutoff), float(dr), _dp(out))
    return out[:k]


def parse_vasp(path: str):
    """io::parse_vasp (reference src/io/vasp_parser.cpp:13-78): (lattice[3,3], frac[N,3], atom_types[N])."""
    lat = np.empty(9)
    cap = 1 << 16
    frac = np.empty((cap, 3)); at = np.empty(cap, np.int32)
    n = lib().ref_parse_vasp(path.encode(), _dp(lat), _dp(frac), at.ctypes.data_as(C.POINTER(C.c_int32)), cap)
    if n < 0:
        raise RuntimeError(lib().ref_last_error().decode())
    return lat.reshape(3, 3), frac[:n].copy(), at[:n].copy()


def build_batch(atom_offsets, frac, lattices, atom_types, r_cutoff, max_neighbors, epsilon, rbf_r_cutoff, dr,
                embed_dims=92, threads=0):
    """Structure + NeighborList + CrystalGraph per structure, std::threads over structures; returns
    (n_edges, checksum, seconds, threads)."""
    ao = np.ascontiguousarray(atom_offsets, dtype=np.int64)
    frac = _f64(frac).reshape(-1, 3)
    lat = _f64(lattices).reshape(-1)
    at = None if atom_types is None else np.ascontiguousarray(atom_types, dtype=np.int32)
    res = _BatchResult()
    rc = lib().ref_build_batch(ao.shape[0] - 1, ao.ctypes.data_as(C.POINTER(C.c_int64)), _dp(frac), _dp(lat),
                               None if at is None else at.ctypes.data_as(C.POINTER(C.c_int32)), float(r_cutoff),
                               C.c_uint64(int(max_neighbors)), float(epsilon), float(rbf_r_cutoff), float(dr),
                               int(embed_dims), int(threads), C.byref(res))
    if rc != 0:
        raise RuntimeError("ref_build_batch failed")
    return int(res.n_edges), float(res.checksum), float(res.seconds), int(res.threads)
