"""The C++ facade that keeps the reference's graph-layer class API (defect-gnn-cpp_b200/host/): NeighborList,
gaussian_rbf, CrystalGraph (drop-in signatures, include/graph/*.hpp of the reference) and the new BatchBuilder.
pytest drives it through the C-callable test shim host/test/facade_shim.cpp, which builds the objects the way a
reference caller does (VASPStructure -> crystal::Structure -> NeighborList -> CrystalGraph)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "defect-gnn-cpp_b200")
UNC = 2**64 - 1


@pytest.fixture(scope="session")
def facade(pkg):
    pkg.build_library()
    pkg.build_host_facade()
    F = C.CDLL(os.path.join(PKG, "libdgnn_facade_test.so"))
    dp, i64p, i32p = C.POINTER(C.c_double), C.POINTER(C.c_int64), C.POINTER(C.c_int32)
    F.facade_last_error.restype = C.c_char_p
    F.facade_positions.argtypes = [C.c_int64, dp, dp, dp]
    F.facade_neighbor_list.argtypes = [C.c_int64, dp, dp, C.c_double, C.c_uint64, C.c_double, C.c_int, C.c_int64, i64p,
                                       i32p, dp, dp, C.POINTER(C.c_int8), i64p, i64p, i64p]
    F.facade_crystal_graph.argtypes = [C.c_int64, dp, dp, i32p, C.c_double, C.c_uint64, C.c_double, C.c_double, C.c_int,
                                       C.c_int, i32p, C.c_int64, i32p, dp, dp, i64p, i32p, i64p]
    F.facade_gaussian_rbf.argtypes = [C.c_double, C.c_double, C.c_double, dp, i32p]
    F.facade_build_batch.argtypes = [C.c_int64, i64p, dp, dp, C.c_double, C.c_uint64, C.c_double, C.c_double, C.c_int64,
                                     i64p, i64p, i32p, i32p, dp, C.POINTER(C.c_float), i64p, i32p]
    return F


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def positions(F, frac, L):
    n = len(frac)
    xyz = np.zeros((n, 3))
    assert F.facade_positions(n, _p(frac, C.c_double), _p(L, C.c_double), _p(xyz, C.c_double)) == 0
    return xyz


def neighbor_list(F, frac, L, rc=10.0, cap=20, eps=1e-10, defaults=False, ecap=200_000):
    n = len(frac)
    rp = np.zeros(n + 1, np.int64)
    dst = np.zeros(ecap, np.int32)
    dist = np.zeros(ecap)
    disp = np.zeros((ecap, 3))
    img = np.zeros((ecap, 3), np.int8)
    ne, near, ties = C.c_int64(), C.c_int64(), C.c_int64()
    rc_ = F.facade_neighbor_list(n, _p(frac, C.c_double), _p(L, C.c_double), rc, C.c_uint64(cap), eps, int(defaults),
                                 ecap, _p(rp, C.c_int64), _p(dst, C.c_int32), _p(dist, C.c_double),
                                 _p(disp, C.c_double), _p(img, C.c_int8), C.byref(ne), C.byref(near), C.byref(ties))
    return rc_, rp, dst[:ne.value], dist[:ne.value], disp[:ne.value], img[:ne.value], near.value, ties.value


# ---- CPU: builds, exports the reference's symbols, input types behave, no silent fallback -----------------------------
def test_facade_exports_reference_api(facade):
    out = subprocess.run(["nm", "-DC", os.path.join(PKG, "libdefect_gnn_graph.so")], capture_output=True, text=True).stdout
    for sym in ("defect_gnn::graph::NeighborList::NeighborList(defect_gnn::crystal::Structure const&, double, unsigned long, double)",
                "defect_gnn::graph::NeighborList::neighbors(unsigned long) const",
                "defect_gnn::graph::gaussian_rbf(double, double, double)",
                "defect_gnn::graph::CrystalGraph::CrystalGraph(",
                "defect_gnn::graph::CrystalGraph::edge_index() const",
                "defect_gnn::graph::CrystalGraph::edge_attr() const",
                "defect_gnn::graph::CrystalGraph::node_features() const",
                "defect_gnn::graph::CrystalGraph::add_topo_features(",
                "defect_gnn::graph::BatchBuilder::build(",
                "defect_gnn::crystal::Structure::Structure(defect_gnn::io::VASPStructure const&)",
                "defect_gnn::io::parse_vasp("):
        assert sym in out, sym


def test_structure_positions_match_oracle(facade, oracle, c1_structures):
    for nm, (x, y, z, L, _, frac) in c1_structures.items():
        xyz = positions(facade, np.ascontiguousarray(frac), np.ascontiguousarray(L))
        ox, oy, oz = oracle.frac_to_cart(frac, L)
        assert np.array_equal(xyz[:, 0], ox) and np.array_equal(xyz[:, 1], oy) and np.array_equal(xyz[:, 2], oz), nm
        assert np.array_equal(xyz[:, 0], x)


def test_facade_fails_loudly_without_gpu(facade, c1_structures):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    _, _, _, L, _, frac = next(iter(c1_structures.values()))
    rc_, *_ = neighbor_list(facade, np.ascontiguousarray(frac), np.ascontiguousarray(L))
    assert rc_ == -22 and b"no CPU fallback" in facade.facade_last_error()


def test_poscar_parser_roundtrip(tmp_path, facade, c1_structures):
    """io::parse_vasp + main_pipeline exist and parse a POSCAR written from a fixture (CPU part: up to the GPU call)."""
    nm, (x, y, z, L, species, frac) = next(iter(c1_structures.items()))
    path = tmp_path / "t.vasp"
    counts = np.bincount(species)
    with open(path, "w") as f:
        f.write("fixture\n1.0\n")
        for r in L:
            f.write(" ".join(repr(float(v)) for v in r) + "\n")
        f.write(" ".join(f"E{i}" for i in range(len(counts))) + "\n" + " ".join(str(int(c)) for c in counts) + "\nDirect\n")
        for p in frac:
            f.write(" ".join(repr(float(v)) for v in p) + "\n")
    r = subprocess.run([os.path.join(PKG, "main_pipeline"), str(path)], capture_output=True, text=True)
    import torch
    if torch.cuda.is_available():
        assert r.returncode == 0 and f"{len(frac)} atoms, {20 * len(frac)} edges, edge_attr {20 * len(frac)} x 100" in r.stdout
    else:
        assert r.returncode == 1 and "no CPU fallback" in r.stderr


def test_bulk_poscar_ingest_matches_per_file_parser(tmp_path, facade, c1_structures):
    """io::load_vasp_batch (multi-threaded POSCAR -> SoA) gives the positions crystal::Structure computes
    from io::parse_vasp's output, file by file: Direct input bit-exact, scaled / Cartesian input to rounding."""
    i32p, i64p, dp = C.POINTER(C.c_int32), C.POINTER(C.c_int64), C.POINTER(C.c_double)
    facade.facade_load_batch.argtypes = [C.c_char_p, C.c_uint, C.c_int64, C.c_int64, i64p, dp, dp, dp, dp, i32p, i64p]
    items = list(c1_structures.items())
    paths, expect = [], []
    for k, (nm, (x, y, z, L, species, frac)) in enumerate(items):
        counts = np.bincount(species)
        scale = 1.0 if k % 3 else 1.5
        cart = k % 4 == 3
        path = tmp_path / f"{nm}.vasp"
        with open(path, "w") as f:
            f.write(f"fixture {nm}\n{scale!r}\n")
            for r in L / scale:
                f.write("  ".join(repr(float(v)) for v in r) + "\n")
            f.write(" ".join(f"E{i}" for i in range(len(counts))) + "\n" + "  ".join(str(int(c)) for c in counts) + "\n")
            if k % 2:
                f.write("Selective dynamics\n")
            f.write("Cartesian\n" if cart else "Direct\n")
            pts = np.stack([x, y, z], axis=1) if cart else frac   # (the reference does not scale Cartesian coordinates)
            for p_ in pts:
                f.write(" ".join(repr(float(v)) for v in p_) + (" T T T" if k % 2 else "") + "\n")
        paths.append(str(path))
        expect.append((scale, cart))
    n_atoms = sum(len(it[1][0]) for it in items)
    for threads in (1, 4):
        offs = np.zeros(len(items) + 1, np.int64)
        X, Y, Z = np.zeros(n_atoms), np.zeros(n_atoms), np.zeros(n_atoms)
        Ls = np.zeros((len(items), 9))
        el = np.zeros(n_atoms, np.int32)
        ns = C.c_int64()
        rc_ = facade.facade_load_batch("\n".join(paths).encode(), threads, n_atoms, len(items), _p(offs, C.c_int64),
                                       _p(X, C.c_double), _p(Y, C.c_double), _p(Z, C.c_double), _p(Ls, C.c_double),
                                       _p(el, C.c_int32), C.byref(ns))
        assert rc_ == 0, facade.facade_last_error()
        assert ns.value == len(items) and offs[-1] == n_atoms
        for k, (nm, (x, y, z, L, species, frac)) in enumerate(items):
            a0, a1 = int(offs[k]), int(offs[k + 1])
            assert a1 - a0 == len(x)
            scale, cart = expect[k]
            assert np.array_equal(el[a0:a1], species)
            got = np.stack([X[a0:a1], Y[a0:a1], Z[a0:a1]], axis=1)
            if scale == 1.0 and not cart:
                assert np.array_equal(Ls[k].reshape(3, 3), L)
                assert np.array_equal(got, positions(facade, np.ascontiguousarray(frac), np.ascontiguousarray(L)))
            else:
                assert np.allclose(Ls[k].reshape(3, 3), L, rtol=1e-15, atol=0)
                assert np.allclose(got, np.stack([x, y, z], axis=1), rtol=0, atol=1e-11)
    # a missing file and a truncated file fail with the parser's error text
    ns = C.c_int64()
    bad = tmp_path / "bad.vasp"
    bad.write_text("x\n1.0\n1 0 0\n")
    for pth in (str(tmp_path / "nope.vasp"), str(bad)):
        rc_ = facade.facade_load_batch(pth.encode(), 2, n_atoms, 1, _p(offs, C.c_int64), _p(X, C.c_double),
                                       _p(Y, C.c_double), _p(Z, C.c_double), _p(Ls, C.c_double), _p(el, C.c_int32),
                                       C.byref(ns))
        assert rc_ != 0 and b"parse_vasp" in facade.facade_last_error()


def test_parsers_keep_the_reference_semantics(tmp_path, facade, c1_structures):
    """io::parse_vasp and io::load_vasp_batch against the reference's own parser + Structure (oracle/_ref, compiled
    verbatim from src/io/vasp_parser.cpp and src/crystal/structure.cpp): Direct iff the mode line STARTS with d / D,
    Cartesian coordinates converted with the scaled lattice's inverse and not scaled themselves -- bit for bit."""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import ref_py as R
    if not R.available():
        pytest.skip("oracle/_ref not built")
    i32p, i64p, dp = C.POINTER(C.c_int32), C.POINTER(C.c_int64), C.POINTER(C.c_double)
    facade.facade_load_batch.argtypes = [C.c_char_p, C.c_uint, C.c_int64, C.c_int64, i64p, dp, dp, dp, dp, i32p, i64p]
    facade.facade_parse_vasp.argtypes = [C.c_char_p, dp, dp, i32p, C.c_int64, i64p]
    items = list(c1_structures.items())
    cases = []
    for k, mode_line in enumerate(["Direct", "direct", "Cartesian", "cartesian", "Kartesisch", " Direct", "", "Direct"]):
        nm, (x, y, z, L, species, frac) = items[k % len(items)]
        n = len(x) - (k % 2)                      # odd and even atom counts (Eigen's packet boundaries differ)
        counts = np.bincount(species[:n])
        scale = (1.0, 1.5, 0.7)[k % 3]
        as_direct = mode_line[:1] in ("d", "D")
        path = tmp_path / f"case{k}.vasp"
        with open(path, "w") as f:
            f.write(f"case {k}\n{scale!r}\n")
            for r in L / scale:
                f.write("  ".join(repr(float(v)) for v in r) + "\n")
            f.write(" ".join(f"E{i}" for i in range(len(counts))) + "\n" + "  ".join(str(int(c)) for c in counts) + "\n")
            f.write(mode_line + "\n")
            pts = frac[:n] if as_direct else np.stack([x, y, z], axis=1)[:n]
            for p_ in pts:
                f.write(" ".join(repr(float(v)) for v in p_) + "\n")
        cases.append((str(path), n))
    for path, n in cases:
        Lr, fr, tr = R.parse_vasp(path)
        st = R.Structure(fr, Lr, tr, 3.0, 1)
        pos_ref = st.neighbor_list().pos
        st.close()
        # per-file parser
        Lf, ff, tf, nn = np.zeros(9), np.zeros((n, 3)), np.zeros(n, np.int32), C.c_int64()
        assert facade.facade_parse_vasp(path.encode(), _p(Lf, C.c_double), _p(ff, C.c_double), _p(tf, C.c_int32), n,
                                        C.byref(nn)) == 0, facade.facade_last_error()
        assert nn.value == n and np.array_equal(tf, tr)
        assert np.array_equal(Lf.reshape(3, 3).view(np.int64), Lr.view(np.int64)), path
        assert np.array_equal(ff.view(np.int64), fr.view(np.int64)), path
        # bulk ingest: positions as crystal::Structure computes them
        offs, X, Y, Z = np.zeros(2, np.int64), np.zeros(n), np.zeros(n), np.zeros(n)
        Ls, el, ns = np.zeros(9), np.zeros(n, np.int32), C.c_int64()
        assert facade.facade_load_batch(path.encode(), 1, n, 1, _p(offs, C.c_int64), _p(X, C.c_double), _p(Y, C.c_double),
                                        _p(Z, C.c_double), _p(Ls, C.c_double), _p(el, C.c_int32), C.byref(ns)) == 0
        assert np.array_equal(Ls.view(np.int64), Lr.reshape(9).view(np.int64)), path
        assert np.array_equal(np.stack([X, Y, Z], axis=1).view(np.int64), pos_ref.view(np.int64)), path


@pytest.mark.gpu
def test_concurrent_constructions_have_their_own_context(facade, oracle, c1_structures):
    """Every calling thread owns a dgnn_ctx (no global mutex): NeighborList + CrystalGraph built on 6 threads at once give
    the same graphs as one thread, and graph::set_device selects the device per thread."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    i64p, dp, ip = C.POINTER(C.c_int64), C.POINTER(C.c_double), C.POINTER(C.c_int)
    facade.facade_concurrent.argtypes = [C.c_int64, dp, dp, C.c_int, C.c_int, C.c_int, i64p, dp, ip]
    x, y, z, L, species, frac = c1_structures["741_1"]
    fr, Lc = np.ascontiguousarray(frac), np.ascontiguousarray(L)
    one_e, one_c, one_r = np.zeros(1, np.int64), np.zeros(1), np.zeros(1, np.int32)
    assert facade.facade_concurrent(len(x), _p(fr, C.c_double), _p(Lc, C.c_double), 1, 3, 0, _p(one_e, C.c_int64),
                                    _p(one_c, C.c_double), _p(one_r, C.c_int)) == 0, facade.facade_last_error()
    assert one_e[0] == 3 * 20 * len(x)
    T = 6
    e, c, r = np.zeros(T, np.int64), np.zeros(T), np.zeros(T, np.int32)
    assert facade.facade_concurrent(len(x), _p(fr, C.c_double), _p(Lc, C.c_double), T, 3, 0, _p(e, C.c_int64),
                                    _p(c, C.c_double), _p(r, C.c_int)) == 0, facade.facade_last_error()
    assert np.all(e == one_e[0]) and np.all(c == one_c[0]) and np.all(r == 3)
    # a device that does not exist is reported, not silently replaced
    assert facade.facade_concurrent(len(x), _p(fr, C.c_double), _p(Lc, C.c_double), 1, 1, 63, _p(one_e, C.c_int64),
                                    _p(one_c, C.c_double), _p(one_r, C.c_int)) != 0
    assert b"device 63" in facade.facade_last_error()


@pytest.mark.gpu
def test_pinned_bulk_ingest_feeds_the_build(tmp_path, facade, oracle, c1_structures):
    """io::load_vasp_batch_pinned: coordinates parsed straight into page-locked SoA buffers, which
    BatchBuilder::build uploads as they are -- same arrays as the pageable ingest, same graphs as the oracle."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    i32p, i64p, dp = C.POINTER(C.c_int32), C.POINTER(C.c_int64), C.POINTER(C.c_double)
    facade.facade_load_batch.argtypes = [C.c_char_p, C.c_uint, C.c_int64, C.c_int64, i64p, dp, dp, dp, dp, i32p, i64p]
    facade.facade_load_batch_pinned_build.argtypes = [C.c_char_p, C.c_uint, C.c_int64, C.c_int64, i64p, dp, dp, dp, dp,
                                                      i32p, i64p, C.c_double, C.c_uint64, i64p]
    items = list(c1_structures.items())
    paths = []
    for nm, (x, y, z, L, species, frac) in items:
        counts = np.bincount(species)
        path = tmp_path / f"{nm}.vasp"
        with open(path, "w") as f:
            f.write(f"fixture {nm}\n1.0\n")
            for r in L:
                f.write("  ".join(repr(float(v)) for v in r) + "\n")
            f.write(" ".join(f"E{i}" for i in range(len(counts))) + "\n" + "  ".join(str(int(c)) for c in counts) + "\nDirect\n")
            for p_ in frac:
                f.write(" ".join(repr(float(v)) for v in p_) + "\n")
        paths.append(str(path))
    n_atoms = sum(len(it[1][0]) for it in items)

    def bufs():
        return (np.zeros(len(items) + 1, np.int64), np.zeros(n_atoms), np.zeros(n_atoms), np.zeros(n_atoms),
                np.zeros((len(items), 9)), np.zeros(n_atoms, np.int32), C.c_int64())
    o1, X1, Y1, Z1, L1, e1, n1 = bufs()
    assert facade.facade_load_batch("\n".join(paths).encode(), 3, n_atoms, len(items), _p(o1, C.c_int64), _p(X1, C.c_double),
                                    _p(Y1, C.c_double), _p(Z1, C.c_double), _p(L1, C.c_double), _p(e1, C.c_int32),
                                    C.byref(n1)) == 0
    for threads in (1, 5):
        o2, X2, Y2, Z2, L2, e2, n2 = bufs()
        ne = C.c_int64()
        rc_ = facade.facade_load_batch_pinned_build("\n".join(paths).encode(), threads, n_atoms, len(items),
                                                    _p(o2, C.c_int64), _p(X2, C.c_double), _p(Y2, C.c_double),
                                                    _p(Z2, C.c_double), _p(L2, C.c_double), _p(e2, C.c_int32), C.byref(n2),
                                                    8.0, C.c_uint64(2**64 - 1), C.byref(ne))
        assert rc_ == 0, facade.facade_last_error()
        for a, b in ((o1, o2), (X1, X2), (Y1, Y2), (Z1, Z2), (L1, L2), (e1, e2)):
            assert np.array_equal(a, b)
        want = int(oracle.batch_edge_counts(o1, X1, Y1, Z1, L1, 8.0).sum())
        assert ne.value == want


# ---- GPU: reference-shaped API against the oracle -------------------------------------------------------------------
@pytest.mark.gpu
def test_neighbor_list_facade_vs_oracle(facade, oracle, c1_structures):
    for nm, (x, y, z, L, _, frac) in list(c1_structures.items())[:4]:
        frac, L = np.ascontiguousarray(frac), np.ascontiguousarray(L)
        for rc, cap, defaults in ((10.0, 20, True), (5.0, 12, False), (8.0, UNC, False)):
            rc_, rp, dst, dist, disp, img, near, ties = neighbor_list(facade, frac, L, rc, cap, defaults=defaults)
            assert rc_ == 0, facade.facade_last_error()
            o = oracle.neighbor_list(x, y, z, L, rc, cap, 1e-10, oracle.ALGO_CELLLIST)
            assert np.array_equal(rp, o.row_ptr) and np.array_equal(dst, o.dst)
            assert np.array_equal(dist.view(np.int64), o.dist.view(np.int64))
            assert np.array_equal(img, o.image[:, :3])
            assert np.array_equal(disp.view(np.int64), o.disp.view(np.int64)), "displacement must be bit-identical"
            assert near == o.n_near_cutoff and ties == o.n_tie_rows


@pytest.mark.gpu
def test_crystal_graph_facade_vs_oracle(facade, oracle, c1_structures):
    nm, (x, y, z, L, species, frac) = next(iter(c1_structures.items()))
    frac, L = np.ascontiguousarray(frac), np.ascontiguousarray(L)
    types = np.ascontiguousarray(species, dtype=np.int32)
    n, dims, ecap = len(frac), 92, 20 * len(frac)
    keys = np.unique(types).astype(np.int32)
    ei = np.zeros(2 * ecap, np.int32)
    ea = np.zeros(ecap * 100)
    nf = np.zeros(n * dims)
    ne, K, nn = C.c_int64(), C.c_int32(), C.c_int64()
    rc_ = facade.facade_crystal_graph(n, _p(frac, C.c_double), _p(L, C.c_double), _p(types, C.c_int32), 10.0,
                                      C.c_uint64(20), 10.0, 0.1, dims, len(keys), _p(keys, C.c_int32), ecap,
                                      _p(ei, C.c_int32), _p(ea, C.c_double), _p(nf, C.c_double), C.byref(ne),
                                      C.byref(K), C.byref(nn))
    assert rc_ == 0, facade.facade_last_error()
    E = ne.value
    assert E == 20 * n and K.value == 100 and nn.value == n
    o = oracle.neighbor_list(x, y, z, L, 10.0, 20, 1e-10, oracle.ALGO_CELLLIST)
    src = np.repeat(np.arange(n), np.diff(o.row_ptr))
    # Eigen::MatrixXi 2 x E column-major: memory is [src0, dst0, src1, dst1, ...]  (crystal_graph.cpp:35-36)
    assert np.array_equal(ei[0:2 * E:2], src) and np.array_equal(ei[1:2 * E:2], o.dst)
    ref = oracle.edge_attr(o.dist, 10.0, 0.1)                       # E x K
    got = ea[:E * 100].reshape(100, E).T                            # column-major E x K
    assert np.max(np.abs(got - ref) / ref) <= 1e-12
    want_nf = (types[:, None] + 0.25 * np.arange(dims)[None, :])    # N x dims, stored column-major
    assert np.array_equal(nf.reshape(dims, n).T, want_nf)
    # a missing embedding key throws std::out_of_range like the reference's map::at (crystal_graph.cpp:20)
    rc_ = facade.facade_crystal_graph(n, _p(frac, C.c_double), _p(L, C.c_double), _p(types, C.c_int32), 10.0,
                                      C.c_uint64(20), 10.0, 0.1, dims, len(keys) - 1, _p(keys, C.c_int32), ecap,
                                      _p(ei, C.c_int32), _p(ea, C.c_double), _p(nf, C.c_double), C.byref(ne),
                                      C.byref(K), C.byref(nn))
    assert rc_ == -20
    # gaussian_rbf(distance, r_cutoff, dr) one distance at a time
    out = np.zeros(128)
    k = C.c_int32()
    for d, rc, dr in ((0.0, 10.0, 0.1), (4.0, 8.0, 0.2), (7.999999, 8.0, 0.2)):
        assert facade.facade_gaussian_rbf(d, rc, dr, _p(out, C.c_double), C.byref(k)) == 0
        ref = oracle.gaussian_rbf(d, rc, dr)
        assert k.value == len(ref) and np.max(np.abs(out[:k.value] - ref) / ref) <= 1e-12


@pytest.mark.gpu
def test_batch_builder_facade(facade, oracle, c1_structures):
    items = list(c1_structures.values())[:3]
    fr = np.ascontiguousarray(np.concatenate([it[5] for it in items]))
    Ls = np.ascontiguousarray(np.stack([it[3] for it in items]))
    offs = np.zeros(len(items) + 1, np.int64)
    offs[1:] = np.cumsum([len(it[5]) for it in items])
    ecap = 120_000
    seo = np.zeros(len(items) + 1, np.int64)
    rp = np.zeros(int(offs[-1]) + 1, np.int64)
    dst = np.zeros(ecap, np.int32)
    img = np.zeros(ecap, np.int32)
    dist = np.zeros(ecap)
    rbf = np.zeros((ecap, 40), np.float32)
    ne, K = C.c_int64(), C.c_int32()
    rc_ = facade.facade_build_batch(len(items), _p(offs, C.c_int64), _p(fr, C.c_double), _p(Ls, C.c_double), 8.0,
                                    C.c_uint64(UNC), 8.0, 0.2, ecap, _p(seo, C.c_int64), _p(rp, C.c_int64),
                                    _p(dst, C.c_int32), _p(img, C.c_int32), _p(dist, C.c_double), _p(rbf, C.c_float),
                                    C.byref(ne), C.byref(K))
    assert rc_ == 0, facade.facade_last_error()
    assert K.value == 40
    e0 = 0
    for s, (x, y, z, L, _, frac) in enumerate(items):
        o = oracle.neighbor_list(x, y, z, L, 8.0, oracle.UNCAPPED, 1e-10, oracle.ALGO_CELLLIST)
        assert seo[s] == e0
        e1 = e0 + o.n_edges
        assert np.array_equal(dst[e0:e1], o.dst)
        assert np.array_equal(dist[e0:e1].view(np.int64), o.dist.view(np.int64))
        assert np.array_equal(img[e0:e1].view(np.int8).reshape(-1, 4), o.image)
        ref = oracle.edge_attr(o.dist, 8.0, 0.2)
        assert np.max(np.abs(rbf[e0:e1] - ref) / ref) <= 1e-5
        e0 = e1
    assert seo[-1] == e0 == ne.value


@pytest.mark.gpu
def test_batch_builder_collate_and_edge_views(facade, oracle, c1_structures):
    """BatchBuilder::collate (the planned train::BatchedGraph, device-resident) and ::edge_views (the viewer's flat
    getters) against the numpy restatements of crystal_graph.cpp:15-40 / wasm_bindings.cpp:206-294."""
    i32p, i64p, fp, dp = C.POINTER(C.c_int32), C.POINTER(C.c_int64), C.POINTER(C.c_float), C.POINTER(C.c_double)
    facade.facade_collate_batch.argtypes = [C.c_int64, i64p, dp, dp, i32p, C.c_double, C.c_uint64, C.c_double, C.c_int,
                                            C.c_int, i32p, C.c_int64, fp, i64p, i64p, i32p, i32p, fp, fp, i64p, i32p]
    items = list(c1_structures.values())[:3]
    fr = np.ascontiguousarray(np.concatenate([it[5] for it in items]))
    Ls = np.ascontiguousarray(np.stack([it[3] for it in items]))
    types = np.ascontiguousarray(np.concatenate([it[4] for it in items]).astype(np.int32))
    offs = np.zeros(len(items) + 1, np.int64)
    offs[1:] = np.cumsum([len(it[5]) for it in items])
    n = int(offs[-1])
    ecap, F = 20 * n, 8
    keys = np.array(sorted(set(types.tolist())), np.int32)

    def run(emb_keys):
        x = np.zeros((n, F), np.float32)
        ei = np.zeros((2, ecap), np.int64)
        batch = np.zeros(n, np.int64)
        src, tgt = np.zeros(ecap, np.int32), np.zeros(ecap, np.int32)
        dist, disp = np.zeros(ecap, np.float32), np.zeros(3 * ecap, np.float32)
        ne, ng = C.c_int64(), C.c_int32()
        rc_ = facade.facade_collate_batch(len(items), _p(offs, C.c_int64), _p(fr, C.c_double), _p(Ls, C.c_double),
                                          _p(types, C.c_int32), 10.0, C.c_uint64(20), 0.1, F, len(emb_keys),
                                          _p(emb_keys, C.c_int32), ecap, _p(x, C.c_float), _p(ei, C.c_int64),
                                          _p(batch, C.c_int64), _p(src, C.c_int32), _p(tgt, C.c_int32),
                                          _p(dist, C.c_float), _p(disp, C.c_float), C.byref(ne), C.byref(ng))
        return rc_, x, ei, batch, src, tgt, dist, disp, ne.value, ng.value

    rc_, x, ei, batch, src, tgt, dist, disp, ne, ng = run(keys)
    assert rc_ == 0, facade.facade_last_error()
    assert ne == ecap and ng == len(items)      # capped at 20 neighbours: E = 20 N, and edge_index is [2, E] exactly
    emb = {int(k): k + 0.25 * np.arange(F) for k in keys}
    og = [oracle.neighbor_list(it[0], it[1], it[2], it[3], 10.0, 20, 1e-10, oracle.ALGO_CELLLIST) for it in items]
    Xo, eio, bo = oracle.collate(og, [it[4] for it in items], emb)
    assert np.array_equal(x, Xo) and np.array_equal(ei, eio) and np.array_equal(batch, bo)
    e0 = 0
    for it, o in zip(items, og):
        so, to, do, po = oracle.viewer_arrays(o, oracle.displacements(it[0], it[1], it[2], it[3], o))
        e1 = e0 + o.n_edges
        assert np.array_equal(src[e0:e1], so) and np.array_equal(tgt[e0:e1], to)
        assert np.array_equal(dist[e0:e1], do) and np.array_equal(disp[3 * e0:3 * e1], po)
        e0 = e1
    # a species without an embedding: std::out_of_range, like CrystalGraph's map::at (crystal_graph.cpp:20)
    rc_ = run(keys[:-1])[0]
    assert rc_ == 2


def test_sharded_partition_matches_python_helper(facade, pkg, syn):
    """graph::predicted_edges / partition_structures (C++ host logic of the in-process sharded builder) against the
    Python helpers bench.py's ranks use (defect-gnn-cpp_b200/sharding.py): every structure assigned exactly once,
    contiguous, balanced by predicted edges."""
    i64p, dp = C.POINTER(C.c_int64), C.POINTER(C.c_double)
    facade.facade_partition.argtypes = [C.c_int64, i64p, dp, C.c_double, C.c_int, i64p, dp]
    sh = pkg.sharding
    none = (np.zeros(0), np.zeros(0), np.zeros(0), np.diag([5.0, 5.0, 5.0]))
    parts = [syn.perovskite_supercell(4, 4, 2, seed=s)[:4] for s in range(5)] + [none, syn.triclinic_cell(3)[:4]] + \
            [syn.perovskite_supercell(2, 2, 2, seed=9)[:4], syn.perovskite_supercell(4, 4, 4, seed=10)[:4]]
    b = syn.concat(parts)
    S = b.n_structs
    offs = np.ascontiguousarray(b.atom_offsets, np.int64)
    L = np.ascontiguousarray(b.lattices.reshape(S, 9))
    w_ref = sh.predicted_edges(offs, b.lattices, 8.0)
    for n in (1, 2, 3, 8, 16):
        bounds = np.zeros(n + 1, np.int64)
        w = np.zeros(S)
        assert facade.facade_partition(S, _p(offs, C.c_int64), _p(L, C.c_double), 8.0, n, _p(bounds, C.c_int64),
                                       _p(w, C.c_double)) == 0, facade.facade_last_error()
        assert np.allclose(w, w_ref, rtol=1e-12)
        assert bounds[0] == 0 and bounds[-1] == S and np.all(np.diff(bounds) >= 0)
        ref = sh.partition_structures(w_ref, n)
        loads = [w_ref[a:c].sum() for a, c in zip(bounds[:-1], bounds[1:])]
        loads_ref = [w_ref[a:c].sum() for a, c in ref]
        assert max(loads) <= max(loads_ref) * (1 + 1e-9) + 1e-9   # never worse balanced than the Python partition


@pytest.mark.gpu
def test_sharded_builder_concatenation_equals_single_build(facade, oracle, c1_structures):
    """graph::ShardedBuilder over two contexts (here: the same device twice; on a multi-GPU box one per device):
    the shards' edge lists concatenated equal the oracle's, global edge offsets are the scan of per-shard counts."""
    i32p, i64p, dp = C.POINTER(C.c_int32), C.POINTER(C.c_int64), C.POINTER(C.c_double)
    facade.facade_build_sharded.argtypes = [C.c_int64, i64p, dp, dp, C.c_double, C.c_uint64, C.c_double, C.c_int, i32p,
                                            C.c_int64, i64p, i64p, i32p, dp, i64p]
    import torch
    items = list(c1_structures.values())[:5]
    fr = np.ascontiguousarray(np.concatenate([it[5] for it in items]))
    Ls = np.ascontiguousarray(np.stack([it[3] for it in items]))
    offs = np.zeros(len(items) + 1, np.int64)
    offs[1:] = np.cumsum([len(it[5]) for it in items])
    ndev = torch.cuda.device_count()
    for n_shards in (1, 2, 3):
        devices = np.array([r % max(ndev, 1) for r in range(n_shards)], np.int32)
        ecap = 20 * int(offs[-1])
        bounds = np.zeros(n_shards + 1, np.int64)
        geo = np.zeros(len(items) + 1, np.int64)
        dst = np.zeros(ecap, np.int32)
        dist = np.zeros(ecap)
        ne = C.c_int64()
        rc_ = facade.facade_build_sharded(len(items), _p(offs, C.c_int64), _p(fr, C.c_double), _p(Ls, C.c_double), 10.0,
                                          C.c_uint64(20), 0.1, n_shards, _p(devices, C.c_int32), ecap,
                                          _p(bounds, C.c_int64), _p(geo, C.c_int64), _p(dst, C.c_int32),
                                          _p(dist, C.c_double), C.byref(ne))
        assert rc_ == 0, facade.facade_last_error()
        assert bounds[0] == 0 and bounds[-1] == len(items) and ne.value == ecap
        e0 = 0
        for s, (x, y, z, L, _, frac) in enumerate(items):
            o = oracle.neighbor_list(x, y, z, L, 10.0, 20, 1e-10, oracle.ALGO_CELLLIST)
            assert geo[s] == e0
            e1 = e0 + o.n_edges
            assert np.array_equal(dst[e0:e1], o.dst)
            assert np.array_equal(dist[e0:e1].view(np.int64), o.dist.view(np.int64))
            e0 = e1
        assert geo[-1] == e0
