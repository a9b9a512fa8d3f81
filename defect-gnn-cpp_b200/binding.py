"""ctypes binding to libdgnn_b200.so -- the C ABI declared in include/dgnn_cabi.h.

This is plumbing for tests / bench / smoke only (Python owns no part of the hot path): it passes plain pointers
through the C ABI exactly like the C++ facade in host/ does.  There is NO CPU fallback: if the shared library or a
CUDA device is missing every call raises.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libdgnn_b200.so")
# (A/B experiments only: DGNN_LIB points at an alternative build of the same sources)
LIB_PATH = os.environ.get("DGNN_LIB", LIB_PATH)

UNCAPPED = 2**64 - 1
INPUT_DEVICE = 0x1
SKIP_RBF = 0x2

EXPORTS = [
    "dgnn_abi_version", "dgnn_strerror", "dgnn_last_error", "dgnn_create", "dgnn_destroy", "dgnn_build",
    "dgnn_copy_to_host", "dgnn_copy_edges_to_host", "dgnn_rbf_expand", "dgnn_rbf_expand_f64",
    "dgnn_last_build_info", "dgnn_stream", "dgnn_set_option",
    "dgnn_collate", "dgnn_copy_collated_to_host", "dgnn_edge_views", "dgnn_neighborhood_clouds",
    "dgnn_last_consumer_info", "dgnn_set_offset_sinks", "dgnn_build_stream", "dgnn_distance_tables", "dgnn_copy_distance_tables_to_host", "dgnn_host_alloc", "dgnn_host_free",
    "dgnn_collated_dlpack", "dgnn_live_exports", "dgnn_dlpack_release", "dgnn_build_serial",
    "dgnn_row_offsets_noalloc",
]


class DgnnBatch(C.Structure):
    _fields_ = [
        ("n_structs", C.c_int64),
        ("atom_offsets", C.c_void_p),
        ("x", C.c_void_p),
        ("y", C.c_void_p),
        ("z", C.c_void_p),
        ("lattice", C.c_void_p),
        ("r_cutoff", C.c_double),
        ("max_neighbors", C.c_uint64),
        ("epsilon", C.c_double),
        ("rbf_r_cutoff", C.c_double),
        ("dr", C.c_double),
        ("flags", C.c_uint32),
    ]


class DgnnGraphs(C.Structure):
    _fields_ = [
        ("n_structs", C.c_int64),
        ("n_atoms", C.c_int64),
        ("n_edges", C.c_int64),
        ("K", C.c_int32),
        ("max_degree", C.c_int32),
        ("row_ptr", C.c_void_p),
        ("dst", C.c_void_p),
        ("image", C.c_void_p),
        ("dist", C.c_void_p),
        ("rbf", C.c_void_p),
        ("struct_edge_offsets", C.c_void_p),
        ("n_near_cutoff", C.c_int64),
        ("n_tie_rows", C.c_int64),
    ]


class DgnnBuildInfo(C.Structure):
    _fields_ = [
        ("h2d_ms", C.c_float), ("bin_ms", C.c_float), ("count_ms", C.c_float), ("fill_ms", C.c_float),
        ("rbf_ms", C.c_float), ("total_ms", C.c_float),
        ("kernel_launches", C.c_int32), ("n_items", C.c_int32), ("cloud_points", C.c_int64),
        ("staged_tile_cap", C.c_int32), ("row_cap", C.c_int32),
        ("n_cell_structs", C.c_int32), ("direct_warps", C.c_int32), ("rbf_fused", C.c_int32), ("hit_cap", C.c_int32),
        ("onepass", C.c_int32), ("onepass_fallback_rows", C.c_int32),
        ("tie_run_mean", C.c_float), ("full_tie_sort", C.c_int32), ("split_big_rows", C.c_int32),
    ]


class DgnnCollated(C.Structure):
    _fields_ = [
        ("n_nodes", C.c_int64), ("n_edges", C.c_int64), ("F", C.c_int32), ("K", C.c_int32),
        ("x", C.c_void_p), ("edge_index", C.c_void_p), ("batch", C.c_void_p), ("edge_attr", C.c_void_p),
    ]


class DgnnConsumerInfo(C.Structure):
    _fields_ = [("device_ms", C.c_float), ("kernel_launches", C.c_int32), ("bytes_read", C.c_int64),
                ("bytes_written", C.c_int64)]


class DgnnDistanceTables(C.Structure):
    _fields_ = [("n_atoms", C.c_int64), ("n_values", C.c_int64), ("values", C.c_void_p), ("table_ptr", C.c_void_p),
                ("mode", C.c_int32)]


class DgnnStreamInfo(C.Structure):
    _fields_ = [("n_chunks", C.c_int64), ("n_structs", C.c_int64), ("n_atoms", C.c_int64), ("n_edges", C.c_int64),
                ("wall_ms", C.c_double), ("build_device_ms", C.c_double), ("h2d_ms", C.c_double),
                ("sink_ms", C.c_double), ("kernel_launches", C.c_int64)]


STREAM_SOURCE = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int64, C.POINTER(DgnnBatch))
STREAM_SINK = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int64, C.POINTER(DgnnGraphs), C.c_void_p)


# ---- DLPack capsules around DLManagedTensor* produced by dgnn_collated_dlpack -----------------------------------------
_CAPSULE_NAME = b"dltensor"
_capsule_destructor_t = C.CFUNCTYPE(None, C.c_void_p)


def _make_capsule_destructor():
    api = C.pythonapi
    api.PyCapsule_IsValid.argtypes = [C.c_void_p, C.c_char_p]
    api.PyCapsule_IsValid.restype = C.c_int
    api.PyCapsule_GetPointer.argtypes = [C.c_void_p, C.c_char_p]
    api.PyCapsule_GetPointer.restype = C.c_void_p

    def _destroy(capsule):
        # a consumer renames the capsule to "used_dltensor" and takes over the deleter; an unconsumed one is ours
        if api.PyCapsule_IsValid(capsule, _CAPSULE_NAME):
            load_library().dgnn_dlpack_release(api.PyCapsule_GetPointer(capsule, _CAPSULE_NAME))
    return _capsule_destructor_t(_destroy)


_capsule_destructor = None


class DLPackArray:
    """One array of a collated batch as a DLPack producer (`__dlpack__` / `__dlpack_device__`): zero-copy into
    torch.from_dlpack, cupy.from_dlpack ...  Each `__dlpack__` call exports a tensor that borrows the builder's arena;
    until the consumer drops it, build() / collate() on that builder raise DGNN_ERR_BUSY instead of overwriting it."""

    def __init__(self, builder, collated, which: int):
        self._gb, self._c, self._which = builder, collated, which

    def __dlpack_device__(self):
        return (2, int(self._gb.device))  # kDLCUDA

    def __dlpack__(self, stream=None, **_kw):
        global _capsule_destructor
        gb = self._gb
        mt = C.c_void_p()
        gb._check(gb.lib.dgnn_collated_dlpack(gb.h, C.byref(self._c), self._which, C.byref(mt)))
        if _capsule_destructor is None:
            _capsule_destructor = _make_capsule_destructor()
        new = C.pythonapi.PyCapsule_New
        new.argtypes = [C.c_void_p, C.c_char_p, _capsule_destructor_t]
        new.restype = C.py_object
        return new(mt, _CAPSULE_NAME, _capsule_destructor)


class DgnnError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"dgnn error {code}: {msg}")
        self.code = code


_lib = None


def load_library():
    """Load libdgnn_b200.so; raises (never falls back) if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FileNotFoundError(
                f"{LIB_PATH} is missing: build it with __graft_entry__.build() (nvcc, sm_100a). "
                "There is no CPU fallback.")
        L = C.CDLL(LIB_PATH)
        L.dgnn_abi_version.restype = C.c_int
        L.dgnn_strerror.argtypes = [C.c_int]
        L.dgnn_strerror.restype = C.c_char_p
        L.dgnn_last_error.argtypes = [C.c_void_p]
        L.dgnn_last_error.restype = C.c_char_p
        L.dgnn_create.argtypes = [C.c_int, C.c_size_t, C.POINTER(C.c_void_p)]
        L.dgnn_create.restype = C.c_int
        L.dgnn_destroy.argtypes = [C.c_void_p]
        L.dgnn_destroy.restype = None
        L.dgnn_build.argtypes = [C.c_void_p, C.POINTER(DgnnBatch), C.POINTER(DgnnGraphs)]
        L.dgnn_build.restype = C.c_int
        L.dgnn_copy_to_host.argtypes = [C.c_void_p, C.POINTER(DgnnGraphs)] + [C.c_void_p] * 6
        L.dgnn_copy_to_host.restype = C.c_int
        L.dgnn_copy_edges_to_host.argtypes = [C.c_void_p, C.POINTER(DgnnGraphs), C.c_int64, C.c_int64] + [C.c_void_p] * 4
        L.dgnn_copy_edges_to_host.restype = C.c_int
        L.dgnn_rbf_expand.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_double, C.c_double, C.c_void_p,
                                      C.POINTER(C.c_int32)]
        L.dgnn_rbf_expand.restype = C.c_int
        L.dgnn_rbf_expand_f64.argtypes = L.dgnn_rbf_expand.argtypes
        L.dgnn_rbf_expand_f64.restype = C.c_int
        L.dgnn_last_build_info.argtypes = [C.c_void_p, C.POINTER(DgnnBuildInfo)]
        L.dgnn_last_build_info.restype = C.c_int
        L.dgnn_stream.argtypes = [C.c_void_p]
        L.dgnn_stream.restype = C.c_void_p
        L.dgnn_set_option.argtypes = [C.c_void_p, C.c_char_p, C.c_double]
        L.dgnn_set_option.restype = C.c_int
        L.dgnn_collate.argtypes = [C.c_void_p, C.POINTER(DgnnGraphs), C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32,
                                   C.c_int32, C.c_uint32, C.POINTER(DgnnCollated)]
        L.dgnn_collate.restype = C.c_int
        L.dgnn_copy_collated_to_host.argtypes = [C.c_void_p, C.POINTER(DgnnCollated)] + [C.c_void_p] * 3
        L.dgnn_copy_collated_to_host.restype = C.c_int
        L.dgnn_edge_views.argtypes = [C.c_void_p, C.POINTER(DgnnGraphs)] + [C.c_void_p] * 4
        L.dgnn_edge_views.restype = C.c_int
        L.dgnn_neighborhood_clouds.argtypes = [C.c_void_p, C.POINTER(DgnnGraphs), C.c_int64, C.c_int64, C.c_void_p,
                                               C.c_void_p]
        L.dgnn_neighborhood_clouds.restype = C.c_int
        L.dgnn_set_offset_sinks.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.dgnn_set_offset_sinks.restype = C.c_int
        L.dgnn_last_consumer_info.argtypes = [C.c_void_p, C.POINTER(DgnnConsumerInfo)]
        L.dgnn_last_consumer_info.restype = C.c_int
        L.dgnn_build_stream.argtypes = [C.c_void_p, STREAM_SOURCE, C.c_void_p, STREAM_SINK, C.c_void_p,
                                        C.POINTER(DgnnStreamInfo)]
        L.dgnn_build_stream.restype = C.c_int
        L.dgnn_distance_tables.argtypes = [C.c_void_p, C.POINTER(DgnnGraphs), C.c_int64, C.c_int64, C.c_int32,
                                           C.POINTER(DgnnDistanceTables)]
        L.dgnn_distance_tables.restype = C.c_int
        L.dgnn_copy_distance_tables_to_host.argtypes = [C.c_void_p, C.POINTER(DgnnDistanceTables), C.c_void_p, C.c_void_p]
        L.dgnn_copy_distance_tables_to_host.restype = C.c_int
        L.dgnn_collated_dlpack.argtypes = [C.c_void_p, C.POINTER(DgnnCollated), C.c_int32, C.POINTER(C.c_void_p)]
        L.dgnn_collated_dlpack.restype = C.c_int
        L.dgnn_live_exports.argtypes = [C.c_void_p]
        L.dgnn_live_exports.restype = C.c_int
        L.dgnn_dlpack_release.argtypes = [C.c_void_p]
        L.dgnn_dlpack_release.restype = None
        L.dgnn_row_offsets_noalloc.argtypes = [C.c_int, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_char_p,
                                               C.c_size_t]
        L.dgnn_row_offsets_noalloc.restype = C.c_int
        _lib = L
    return _lib


NOALLOC_MAX_ROWS = 1 << 24  # DGNN_NOALLOC_MAX_ROWS (include/dgnn_cabi.h)


def row_offsets_noalloc(device: int, degrees_ptr: int, n_rows: int, row_ptr_ptr: int, stream: int | None = None) -> None:
    """dgnn_row_offsets_noalloc over raw DEVICE pointers (int32 degrees[n_rows] -> int64 row_ptr[n_rows + 1]) on
    `stream` (cudaStream_t as int, None = default stream); does not synchronise and allocates nothing.  No counterpart
    in the reference (see include/dgnn_cabi.h)."""
    lib = load_library()
    err = C.create_string_buffer(256)
    rc = lib.dgnn_row_offsets_noalloc(int(device), C.c_void_p(degrees_ptr), int(n_rows), C.c_void_p(row_ptr_ptr),
                                      C.c_void_p(stream), err, len(err))
    if rc != 0:
        raise DgnnError(rc, err.value.decode() or lib.dgnn_strerror(rc).decode())


@dataclass
class HostGraphs:
    """Host copy of a build result (numpy)."""
    n_structs: int
    n_atoms: int
    n_edges: int
    K: int
    max_degree: int
    row_ptr: np.ndarray
    dst: np.ndarray
    image: np.ndarray  # int8 [E,4]
    dist: np.ndarray
    rbf: np.ndarray | None
    struct_edge_offsets: np.ndarray
    n_near_cutoff: int
    n_tie_rows: int


def _ptr(a) -> int | None:
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data
    if hasattr(a, "data_ptr"):  # torch tensor (host pinned or device)
        return a.data_ptr()
    return int(a)


class GraphBuilder:
    """One dgnn_ctx (one CUDA device).  Mirrors the call sequence of the C++ facade."""

    def __init__(self, device: int = 0, arena_bytes: int = 0):
        self.lib = load_library()
        h = C.c_void_p()
        rc = self.lib.dgnn_create(int(device), int(arena_bytes), C.byref(h))
        if rc != 0:
            raise DgnnError(rc, self.lib.dgnn_strerror(rc).decode() + " (dgnn_create; a B200 is required, no CPU fallback)")
        self.h = h
        self.device = device
        self.last = None

    def close(self):
        if getattr(self, "_borrowed", False):
            self.h = None
            return
        if getattr(self, "h", None):
            self.lib.dgnn_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc: int):
        if rc != 0:
            raise DgnnError(rc, self.lib.dgnn_last_error(self.h).decode() or self.lib.dgnn_strerror(rc).decode())

    def set_option(self, name: str, value: float):
        self._check(self.lib.dgnn_set_option(self.h, name.encode(), float(value)))

    @property
    def stream(self) -> int:
        return int(self.lib.dgnn_stream(self.h) or 0)

    def build(self, atom_offsets, x, y, z, lattices, r_cutoff=10.0, max_neighbors=20, epsilon=1e-10,
              rbf_r_cutoff=10.0, dr=0.1, device_inputs=False, skip_rbf=False) -> DgnnGraphs:
        """dgnn_build.  Inputs: numpy arrays / pinned torch tensors (host) or device tensors (device_inputs=True)."""
        if isinstance(atom_offsets, np.ndarray):
            atom_offsets = np.ascontiguousarray(atom_offsets, dtype=np.int64)
            x, y, z = (np.ascontiguousarray(a, dtype=np.float64) for a in (x, y, z))
            lattices = np.ascontiguousarray(lattices, dtype=np.float64)
            n_structs = atom_offsets.shape[0] - 1
        else:
            n_structs = int(atom_offsets.shape[0]) - 1
        self._keep = (atom_offsets, x, y, z, lattices)  # keep inputs alive during the call
        b = DgnnBatch(n_structs, _ptr(atom_offsets), _ptr(x), _ptr(y), _ptr(z), _ptr(lattices), float(r_cutoff),
                      C.c_uint64(int(max_neighbors)), float(epsilon), float(rbf_r_cutoff),
                      float(dr if dr is not None else 0.0),
                      (INPUT_DEVICE if device_inputs else 0) | (SKIP_RBF if (skip_rbf or dr is None) else 0))
        g = DgnnGraphs()
        self._check(self.lib.dgnn_build(self.h, C.byref(b), C.byref(g)))
        self.last = g
        return g

    def build_stream(self, source, sink=None, r_cutoff=10.0, max_neighbors=20, epsilon=1e-10, rbf_r_cutoff=10.0,
                     dr=0.1, device_inputs=False) -> DgnnStreamInfo:
        """dgnn_build_stream: `source(c)` returns the chunk's (atom_offsets, x, y, z, lattices) -- numpy / pinned torch
        (host) or device tensors -- or None when the job is finished; `sink(c, graphs, arena)` sees the chunk's DgnnGraphs
        (device pointers, valid until it returns) and a GraphBuilder view of the context that owns them (use IT for
        to_host / collate / distance_tables of this chunk).  Chunk c + 1 is uploaded and built while sink(c) runs."""
        keep = {}
        errs = []

        def _src(_user, c, bp):
            try:
                t = source(int(c))
                if t is None:
                    return 0
                off, x, y, z, lat = t
                if isinstance(off, np.ndarray):
                    off = np.ascontiguousarray(off, dtype=np.int64)
                    x, y, z = (np.ascontiguousarray(a, dtype=np.float64) for a in (x, y, z))
                    lat = np.ascontiguousarray(lat, dtype=np.float64)
                keep[int(c) & 1] = (off, x, y, z, lat)  # alive until the chunk after next is asked for
                b = bp.contents
                b.n_structs = int(off.shape[0]) - 1
                b.atom_offsets, b.x, b.y, b.z, b.lattice = _ptr(off), _ptr(x), _ptr(y), _ptr(z), _ptr(lat)
                b.r_cutoff, b.max_neighbors, b.epsilon = float(r_cutoff), int(max_neighbors), float(epsilon)
                b.rbf_r_cutoff, b.dr = float(rbf_r_cutoff), float(dr if dr is not None else 0.0)
                b.flags = (INPUT_DEVICE if device_inputs else 0) | (SKIP_RBF if dr is None else 0)
                return 1
            except Exception as exc:  # noqa: BLE001 -- must not propagate through the C frames
                errs.append(exc)
                return -1

        def _snk(_user, c, gp, arena):
            try:
                if sink is not None:
                    view = GraphBuilder.__new__(GraphBuilder)   # borrowed handle: never destroyed from here
                    view.lib, view.h, view.device, view.last = self.lib, None, self.device, gp.contents
                    view.__dict__["h"] = C.c_void_p(arena)
                    view._borrowed = True
                    sink(int(c), gp.contents, view)
                    view.__dict__["h"] = None
                return 0
            except Exception as exc:  # noqa: BLE001
                errs.append(exc)
                return 1

        info = DgnnStreamInfo()
        rc = self.lib.dgnn_build_stream(self.h, STREAM_SOURCE(_src), None, STREAM_SINK(_snk), None, C.byref(info))
        if errs:
            raise errs[0]
        self._check(rc)
        return info

    def set_offset_sinks(self, row_ptr=None, struct_edge_offsets=None):
        """dgnn_set_offset_sinks: host (pinned) int64 buffers that receive row_ptr / struct_edge_offsets of every
        following build, copied back while the fill pass runs."""
        self._sinks = (row_ptr, struct_edge_offsets)
        self._check(self.lib.dgnn_set_offset_sinks(self.h, _ptr(row_ptr), _ptr(struct_edge_offsets)))

    def info(self) -> DgnnBuildInfo:
        i = DgnnBuildInfo()
        self._check(self.lib.dgnn_last_build_info(self.h, C.byref(i)))
        return i

    def to_host(self, g: DgnnGraphs | None = None, with_rbf: bool = True) -> HostGraphs:
        g = g or self.last
        e = int(g.n_edges)
        row_ptr = np.empty(g.n_atoms + 1, dtype=np.int64)
        seo = np.empty(g.n_structs + 1, dtype=np.int64)
        dst = np.empty(e, dtype=np.int32)
        image = np.empty(e, dtype=np.int32)
        dist = np.empty(e, dtype=np.float64)
        rbf = np.empty((e, g.K), dtype=np.float32) if (with_rbf and g.K > 0 and g.rbf) else None
        self._check(self.lib.dgnn_copy_to_host(self.h, C.byref(g), _ptr(row_ptr), _ptr(dst), _ptr(image), _ptr(dist),
                                               _ptr(rbf), _ptr(seo)))
        return HostGraphs(int(g.n_structs), int(g.n_atoms), e, int(g.K), int(g.max_degree), row_ptr, dst,
                          image.view(np.int8).reshape(e, 4), dist, rbf, seo, int(g.n_near_cutoff),
                          int(g.n_tie_rows))

    def edges_to_host(self, e0: int, e1: int, g: DgnnGraphs | None = None, with_rbf: bool = True):
        g = g or self.last
        n = e1 - e0
        dst = np.empty(n, dtype=np.int32)
        image = np.empty(n, dtype=np.int32)
        dist = np.empty(n, dtype=np.float64)
        rbf = np.empty((n, g.K), dtype=np.float32) if (with_rbf and g.K > 0 and g.rbf) else None
        self._check(self.lib.dgnn_copy_edges_to_host(self.h, C.byref(g), e0, e1, _ptr(dst), _ptr(image), _ptr(dist),
                                                     _ptr(rbf)))
        return dst, image.view(np.int8).reshape(n, 4), dist, rbf

    def rbf_expand(self, dist, r_cutoff: float, dr: float, out=None, f64_colmajor: bool = False):
        """graph::gaussian_rbf for many distances (edge_features.cpp:7-24).  dist/out: numpy or device tensors."""
        n = int(dist.shape[0])
        k = C.c_int32()
        if isinstance(dist, np.ndarray):
            dist = np.ascontiguousarray(dist, dtype=np.float64)
        kk = int(np.floor(r_cutoff / dr))
        if out is None:
            out = np.empty((kk, n), dtype=np.float64) if f64_colmajor else np.empty((n, kk), dtype=np.float32)
        fn = self.lib.dgnn_rbf_expand_f64 if f64_colmajor else self.lib.dgnn_rbf_expand
        self._check(fn(self.h, _ptr(dist), n, float(r_cutoff), float(dr), _ptr(out), C.byref(k)))
        assert k.value == kk
        if f64_colmajor and isinstance(out, np.ndarray):
            return out.T  # column-major E x K viewed as E x K
        return out

    # ---- consumers of the last build's CSR ----
    def collate(self, node_keys=None, emb_keys=None, emb_table=None, g: DgnnGraphs | None = None,
                device_keys: bool = False) -> DgnnCollated:
        """dgnn_collate: PyG-style batch (x, edge_index, batch, edge_attr) left on the device."""
        g = g or self.last
        n_keys = F = 0
        if emb_keys is not None:
            emb_keys = np.ascontiguousarray(emb_keys, dtype=np.int32)
            emb_table = np.ascontiguousarray(emb_table, dtype=np.float64)
            n_keys, F = emb_table.shape
            if isinstance(node_keys, np.ndarray):
                node_keys = np.ascontiguousarray(node_keys, dtype=np.int32)
        self._keep_c = (node_keys, emb_keys, emb_table)
        out = DgnnCollated()
        self._check(self.lib.dgnn_collate(self.h, C.byref(g), _ptr(node_keys), _ptr(emb_keys), _ptr(emb_table),
                                          int(n_keys), int(F), INPUT_DEVICE if device_keys else 0, C.byref(out)))
        return out

    def collated_dlpack(self, c: DgnnCollated) -> dict:
        """DLPack producers for the collated batch: {"x", "edge_index", "batch", "edge_attr"} (absent arrays left out)."""
        out = {"edge_index": DLPackArray(self, c, 1), "batch": DLPackArray(self, c, 2)}
        if c.x:
            out["x"] = DLPackArray(self, c, 0)
        if c.edge_attr and c.K > 0:
            out["edge_attr"] = DLPackArray(self, c, 3)
        return out

    def live_exports(self) -> int:
        return int(self.lib.dgnn_live_exports(self.h))

    def collated_to_host(self, c: DgnnCollated):
        x = np.empty((c.n_nodes, c.F), dtype=np.float32) if c.x else None
        ei = np.empty((2, c.n_edges), dtype=np.int64)
        batch = np.empty(c.n_nodes, dtype=np.int64)
        self._check(self.lib.dgnn_copy_collated_to_host(self.h, C.byref(c), _ptr(x), _ptr(ei), _ptr(batch)))
        return x, ei, batch

    def edge_views(self, g: DgnnGraphs | None = None, displacements: bool = True):
        """dgnn_edge_views: the viewer's flat arrays (sources, targets, distances, displacements)."""
        g = g or self.last
        e = int(g.n_edges)
        src = np.empty(e, dtype=np.int32)
        tgt = np.empty(e, dtype=np.int32)
        dist = np.empty(e, dtype=np.float32)
        disp = np.empty(3 * e, dtype=np.float32) if displacements else None
        self._check(self.lib.dgnn_edge_views(self.h, C.byref(g), _ptr(src), _ptr(tgt), _ptr(dist), _ptr(disp)))
        return src, tgt, dist, disp

    def neighborhood_clouds(self, a0: int, a1: int, row_ptr: np.ndarray, g: DgnnGraphs | None = None):
        """dgnn_neighborhood_clouds for the atoms [a0, a1); row_ptr (host) sizes the output."""
        g = g or self.last
        n_points = int(row_ptr[a1] - row_ptr[a0]) + (a1 - a0)
        cloud = np.empty((n_points, 3), dtype=np.float64)
        cptr = np.empty(a1 - a0 + 1, dtype=np.int64)
        self._check(self.lib.dgnn_neighborhood_clouds(self.h, C.byref(g), int(a0), int(a1), _ptr(cloud), _ptr(cptr)))
        return cloud, cptr

    def distance_tables(self, a0: int, a1: int, mode: int = 0, g: DgnnGraphs | None = None, to_host: bool = True):
        """dgnn_distance_tables for the atoms [a0, a1): mode 0 = full (deg+1)^2 fp64 tables (column-major), mode 1 = the
        compressed lower triangles in fp32 (Ripser's input).  Returns (values, table_ptr) as numpy arrays, or the
        device-pointer record with to_host=False."""
        g = g or self.last
        t = DgnnDistanceTables()
        self._check(self.lib.dgnn_distance_tables(self.h, C.byref(g), int(a0), int(a1), int(mode), C.byref(t)))
        if not to_host:
            return t
        vals = np.empty(int(t.n_values), dtype=np.float64 if mode == 0 else np.float32)
        tptr = np.empty(int(t.n_atoms) + 1, dtype=np.int64)
        self._check(self.lib.dgnn_copy_distance_tables_to_host(self.h, C.byref(t), _ptr(vals), _ptr(tptr)))
        return vals, tptr

    def consumer_info(self) -> DgnnConsumerInfo:
        i = DgnnConsumerInfo()
        self._check(self.lib.dgnn_last_consumer_info(self.h, C.byref(i)))
        return i
