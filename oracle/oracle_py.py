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


# ---------------------------------------------------------------------------------------------------------------------
# Consumers of the neighbour list -- numpy restatements, test infrastructure only.
# ---------------------------------------------------------------------------------------------------------------------
def crystal_graph_arrays(o: OracleGraph, node_keys, atom_embeddings: dict):
    """CrystalGraph's node / edge-index assembly for ONE structure (src/graph/crystal_graph.cpp:15-40):
    node_features.row(i) = atom_embeddings.at(element_i)  (KeyError where std::map::at throws);
    edge_index(0, e) = i, edge_index(1, e) = neighbor.idx, edges by source atom then row order."""
    n = o.row_ptr.shape[0] - 1
    x = np.stack([np.asarray(atom_embeddings[int(k)], dtype=np.float64) for k in node_keys]) if n else np.zeros((0, 0))
    src = np.repeat(np.arange(n, dtype=np.int64), np.diff(o.row_ptr))
    edge_index = np.stack([src, o.dst.astype(np.int64)])
    return x, edge_index


def collate(graphs, node_keys_per_graph, atom_embeddings: dict):
    """The batch collation the training consumer does (planned train::BatchedGraph, PROJECT_ROADMAP.md:892-904; PyG
    Batch semantics of the Python ancestor): concatenate node features, offset edge indices by the node count of the
    preceding graphs, batch[i] = graph of node i.  dtypes of utilities.py:69-71 (x float32, edge_index int64)."""
    xs, eis, batch = [], [], []
    base = 0
    for gi, (o, keys) in enumerate(zip(graphs, node_keys_per_graph)):
        x, ei = crystal_graph_arrays(o, keys, atom_embeddings)
        n = o.row_ptr.shape[0] - 1
        if n:
            xs.append(x.astype(np.float32))
        eis.append(ei + base)
        batch.append(np.full(n, gi, dtype=np.int64))
        base += n
    F = len(next(iter(atom_embeddings.values()))) if atom_embeddings else 0
    x = np.concatenate(xs) if xs else np.zeros((0, F), dtype=np.float32)
    return x, np.concatenate(eis, axis=1) if eis else np.zeros((2, 0), np.int64), \
        np.concatenate(batch) if batch else np.zeros(0, np.int64)


def displacements(x, y, z, lattice, o: OracleGraph) -> np.ndarray:
    """Neighbor::displacement re-derived from (dst, image): cloud_pos - query_pos with
    cloud_pos = pos_j + ((n_a*row0 + n_b*row1) + n_c*row2)  (src/graph/neighbor_list.cpp:82-87, 51); numpy fp64
    element-wise arithmetic rounds once per operation like the reference's FMA-free build.  The exhaustive and KD-tree
    algorithms of the C++ oracle return the same quantity directly (OracleGraph.disp) -- tests pin one on the other."""
    Lm = _f64(lattice).reshape(3, 3)
    pos = np.stack([_f64(x), _f64(y), _f64(z)], axis=1)
    n = o.row_ptr.shape[0] - 1
    src = np.repeat(np.arange(n), np.diff(o.row_ptr))
    na = o.image[:, 0].astype(np.float64)[:, None]
    nb = o.image[:, 1].astype(np.float64)[:, None]
    nc = o.image[:, 2].astype(np.float64)[:, None]
    off = (na * Lm[0][None, :] + nb * Lm[1][None, :]) + nc * Lm[2][None, :]
    return (pos[o.dst] + off) - pos[src]


def viewer_arrays(o: OracleGraph, disp=None):
    """WasmAPI::get_edge_sources / targets / distances / displacements (src/viz/wasm_bindings.cpp:206-294):
    static_cast<int>(i), static_cast<int>(neighbor.idx), static_cast<float>(distance), static_cast<float>(disp.xyz)."""
    n = o.row_ptr.shape[0] - 1
    src = np.repeat(np.arange(n, dtype=np.int32), np.diff(o.row_ptr))
    d = o.disp if disp is None else disp
    return src, o.dst.astype(np.int32), o.dist.astype(np.float32), d.astype(np.float32).reshape(-1)


def neighborhood_cloud(x, y, z, o: OracleGraph, atom: int, disp=None) -> np.ndarray:
    """compute_atom_betti_features' point cloud (src/topology/betti_features.cpp:65-75): row 0 = the atom's position,
    row 1 + e = position + neighbors[e].displacement (fp64, one rounding per component)."""
    e0, e1 = int(o.row_ptr[atom]), int(o.row_ptr[atom + 1])
    c = np.array([x[atom], y[atom], z[atom]], dtype=np.float64)
    out = np.empty((e1 - e0 + 1, 3), dtype=np.float64)
    out[0] = c
    out[1:] = c[None, :] + (o.disp if disp is None else disp)[e0:e1]
    return out


def distance_table(cloud: np.ndarray) -> np.ndarray:
    """compute_persistence's pairwise-distance matrix (src/topology/ripser_wrapper.cpp:60-66), restated with Eigen 3.4's
    FMA-free operation order: sq_norms = rowwise().squaredNorm() = (x x + y y) + z z; sq_dist = (s_i + s_j) - 2 P P^T with
    the depth-3 product accumulated in order ((x_i x_j + y_i y_j) + z_i z_j, scaled by 2 exactly); cwiseMax(0).cwiseSqrt().
    numpy's element-wise fp64 arithmetic rounds once per operation, like the reference's build without -march."""
    P = _f64(cloud).reshape(-1, 3)
    x, y, z = P[:, 0], P[:, 1], P[:, 2]
    s = (x * x + y * y) + z * z
    dot = (x[:, None] * x[None, :] + y[:, None] * y[None, :]) + z[:, None] * z[None, :]
    d2 = (s[:, None] + s[None, :]) - 2.0 * dot
    return np.sqrt(np.maximum(d2, 0.0))


def ripser_lower_triangle(dist: np.ndarray) -> np.ndarray:
    """compute_persistence_from_distances' `distances` vector (ripser_wrapper.cpp:18-24): for i = 1..N-1, j = 0..i-1,
    static_cast<float>(distance_matrix(i, j))."""
    n = dist.shape[0]
    out = [dist[i, :i].astype(np.float32) for i in range(1, n)]
    return np.concatenate(out) if out else np.zeros(0, dtype=np.float32)
