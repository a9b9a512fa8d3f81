import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import __graft_entry__ as entry  # noqa: E402


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def pkg():
    return entry.load_package()


@pytest.fixture(scope="session")
def oracle():
    return entry.load_oracle()


@pytest.fixture(scope="session")
def syn(pkg):
    return pkg.synthetic


@pytest.fixture(scope="session")
def golden_dir():
    return os.path.join(ROOT, "tests", "golden")


@pytest.fixture(scope="session")
def c1_structures(golden_dir):
    d = np.load(os.path.join(golden_dir, "c1_structures.npz"))
    out = {}
    for nm in d["names"]:
        nm = str(nm)
        c = d[f"{nm}__cart"]
        out[nm] = (np.ascontiguousarray(c[:, 0]), np.ascontiguousarray(c[:, 1]), np.ascontiguousarray(c[:, 2]),
                   d[f"{nm}__lattice"].copy(), d[f"{nm}__species"].copy(), d[f"{nm}__frac"].copy())
    return out


@pytest.fixture(scope="session")
def harness_lib():
    """CPU emulation of the product's cell-list logic built on the product's own geom.h (tests/cpu_harness)."""
    import ctypes as C
    hdir = os.path.join(ROOT, "tests", "cpu_harness")
    so = os.path.join(hdir, "libcell_harness.so")
    src = os.path.join(hdir, "cell_harness.cpp")
    geom = os.path.join(ROOT, "defect-gnn-cpp_b200", "csrc", "geom.h")
    if (not os.path.exists(so)) or os.path.getmtime(so) < max(os.path.getmtime(src), os.path.getmtime(geom)):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-ffp-contract=off", "-shared", "-o", so, src])
    H = C.CDLL(so)
    dp, i64p = C.POINTER(C.c_double), C.POINTER(C.c_int64)
    H.harness_neighbor_list.argtypes = [C.c_int64, dp, dp, dp, dp, C.c_double, C.c_uint64, C.c_double, C.c_double,
                                        C.c_int, C.c_int, C.c_int64, i64p, C.POINTER(C.c_int32),
                                        C.POINTER(C.c_uint32), dp, i64p, i64p, i64p]
    H.harness_neighbor_list.restype = C.c_int
    H.harness_direct_neighbor_list.argtypes = [C.c_int64, dp, dp, dp, dp, C.c_double, C.c_uint64, C.c_double, C.c_int64,
                                               i64p, C.POINTER(C.c_int32), C.POINTER(C.c_uint32), dp, i64p, i64p, i64p]
    H.harness_direct_neighbor_list.restype = C.c_int
    return H


@pytest.fixture(scope="session")
def gpu_builder(pkg):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    gb = pkg.GraphBuilder(0)
    yield gb
    gb.close()


@pytest.fixture(scope="session")
def gpu_builder_cell(pkg):
    """Same library, every structure forced through the cell-list path (direct_max_atoms = 0)."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    gb = pkg.GraphBuilder(0)
    gb.set_option("direct_max_atoms", 0)
    yield gb
    gb.close()


def assert_graph_equal(h, e0, a0, a1, o, what=""):
    """Compare the slice of a HostGraphs belonging to atoms [a0,a1) (edges from e0) with an OracleGraph: bit-exact."""
    rp = h.row_ptr[a0:a1 + 1] - h.row_ptr[a0]
    assert np.array_equal(rp, o.row_ptr), f"{what}: row_ptr differs"
    e1 = e0 + o.n_edges
    assert np.array_equal(h.dst[e0:e1], o.dst), f"{what}: dst differs"
    assert np.array_equal(h.image[e0:e1], o.image), f"{what}: image offsets differ"
    assert np.array_equal(h.dist[e0:e1].view(np.int64), o.dist.view(np.int64)), f"{what}: distances not bit-exact"
    return e1
