"""CPU tests of the product's cell-list logic (geom.h: grid sizing, halo-pruned image cloud, work-item tiles,
stencil rows) through tests/cpu_harness, which emulates the kernels' algorithm sequentially using the same header
the CUDA code includes.  Compared bit-exactly with the oracle."""
import ctypes as C

import numpy as np
import pytest


def run_harness(H, x, y, z, L, rc, cap, eps=1e-10, cf=0.5, tile_target=2048, src_target=512, ecap=400_000):
    n = len(x)
    dp, i64p = C.POINTER(C.c_double), C.POINTER(C.c_int64)
    x, y, z = (np.ascontiguousarray(a, dtype=np.float64) for a in (x, y, z))
    Lf = np.ascontiguousarray(L, dtype=np.float64).reshape(9)
    rp = np.zeros(n + 1, np.int64)
    dst = np.zeros(ecap, np.int32)
    img = np.zeros(ecap, np.uint32)
    dist = np.zeros(ecap)
    ne, nn = C.c_int64(), C.c_int64()
    info = np.zeros(10, np.int64)
    rc_ = H.harness_neighbor_list(n, x.ctypes.data_as(dp), y.ctypes.data_as(dp), z.ctypes.data_as(dp),
                                  Lf.ctypes.data_as(dp), rc, C.c_uint64(cap), eps, cf, tile_target, src_target, ecap,
                                  rp.ctypes.data_as(i64p), dst.ctypes.data_as(C.POINTER(C.c_int32)),
                                  img.ctypes.data_as(C.POINTER(C.c_uint32)), dist.ctypes.data_as(dp), C.byref(ne),
                                  C.byref(nn), info.ctypes.data_as(i64p))
    assert rc_ == 0, f"harness rc={rc_}"
    e = ne.value
    return rp, dst[:e], img[:e].view(np.int8).reshape(e, 4), dist[:e], nn.value, info


def check(H, oracle, x, y, z, L, rc, cap=None, algo=None, **kw):
    cap = oracle.UNCAPPED if cap is None else cap
    g = oracle.neighbor_list(x, y, z, L, rc, cap, kw.get("eps", 1e-10), oracle.ALGO_CELLLIST if algo is None else algo)
    rp, dst, img, dist, near, info = run_harness(H, x, y, z, L, rc, cap, **kw)
    assert np.array_equal(rp, g.row_ptr)
    assert np.array_equal(dst, g.dst)
    assert np.array_equal(img, g.image)
    assert np.array_equal(dist.view(np.int64), g.dist.view(np.int64))
    assert near == g.n_near_cutoff
    return info


def test_c2_grids_and_tilings(harness_lib, oracle, syn):
    x, y, z, L, _ = syn.perovskite_supercell(4, 4, 2, seed=1234)
    for cf in (1.0, 0.5, 0.33, 0.2):
        check(harness_lib, oracle, x, y, z, L, 8.0, cf=cf)
    info = check(harness_lib, oracle, x, y, z, L, 8.0, tile_target=300, src_target=20)
    assert info[6] > 8  # forced block decomposition really happened
    check(harness_lib, oracle, x, y, z, L, 10.0, cap=20)
    check(harness_lib, oracle, x, y, z, L, 5.0, cap=12)


def test_triclinic_and_adversarial(harness_lib, oracle, syn):
    b = syn.c3_batch(40)
    for s in range(b.n_structs):
        check(harness_lib, oracle, *b.structure(s), 8.0, algo=oracle.ALGO_BRUTE)
    adv = syn.adversarial_cells()
    for s in range(adv.n_structs):
        check(harness_lib, oracle, *adv.structure(s), 8.0, algo=oracle.ALGO_BRUTE)
        check(harness_lib, oracle, *adv.structure(s), 3.0, cap=5, algo=oracle.ALGO_BRUTE)


def test_shipped_poscars(harness_lib, oracle, c1_structures):
    for nm, (x, y, z, L, _, _) in c1_structures.items():
        check(harness_lib, oracle, x, y, z, L, 10.0)
        check(harness_lib, oracle, x, y, z, L, 10.0, cap=20)


def test_multi_item_supercell_and_commensurate_cutoff(harness_lib, oracle, syn):
    x, y, z, L, _ = syn.perovskite_supercell(7, 6, 5, seed=77)
    info = check(harness_lib, oracle, x, y, z, L, 8.0, tile_target=600, src_target=64, ecap=400_000)
    assert info[6] > 50
    b = syn.simple_cubic(2.0, 3)  # ideal lattice, neighbours at exactly r_c: decided by rounding, must still match
    check(harness_lib, oracle, *b.structure(0), 4.0, algo=oracle.ALGO_BRUTE)
    check(harness_lib, oracle, *b.structure(0), 2.0 * np.sqrt(2.0), algo=oracle.ALGO_BRUTE)


# ---- direct (small-structure) traversal: fractional image windows ---------------------------------------------------
def run_direct(H, x, y, z, L, rc, cap, eps=1e-10, ecap=400_000):
    n = len(x)
    dp, i64p = C.POINTER(C.c_double), C.POINTER(C.c_int64)
    x, y, z = (np.ascontiguousarray(a, dtype=np.float64) for a in (x, y, z))
    Lf = np.ascontiguousarray(L, dtype=np.float64).reshape(9)
    rp = np.zeros(n + 1, np.int64)
    dst = np.zeros(ecap, np.int32)
    img = np.zeros(ecap, np.uint32)
    dist = np.zeros(ecap)
    ne, nn = C.c_int64(), C.c_int64()
    info = np.zeros(4, np.int64)
    rc_ = H.harness_direct_neighbor_list(n, x.ctypes.data_as(dp), y.ctypes.data_as(dp), z.ctypes.data_as(dp),
                                         Lf.ctypes.data_as(dp), rc, C.c_uint64(cap), eps, ecap,
                                         rp.ctypes.data_as(i64p), dst.ctypes.data_as(C.POINTER(C.c_int32)),
                                         img.ctypes.data_as(C.POINTER(C.c_uint32)), dist.ctypes.data_as(dp),
                                         C.byref(ne), C.byref(nn), info.ctypes.data_as(i64p))
    assert rc_ == 0, f"harness rc={rc_}"
    e = ne.value
    return rp, dst[:e], img[:e].view(np.int8).reshape(e, 4), dist[:e], nn.value, info


def check_direct(H, oracle, x, y, z, L, rc, cap=None, algo=None, eps=1e-10):
    cap = oracle.UNCAPPED if cap is None else cap
    g = oracle.neighbor_list(x, y, z, L, rc, cap, eps, oracle.ALGO_CELLLIST if algo is None else algo)
    rp, dst, img, dist, near, info = run_direct(H, x, y, z, L, rc, cap, eps)
    assert np.array_equal(rp, g.row_ptr)
    assert np.array_equal(dst, g.dst)
    assert np.array_equal(img, g.image)
    assert np.array_equal(dist.view(np.int64), g.dist.view(np.int64))
    assert near == g.n_near_cutoff
    return info


def test_direct_windows_c2_and_poscars(harness_lib, oracle, syn, c1_structures):
    x, y, z, L, _ = syn.perovskite_supercell(4, 4, 2, seed=1234)
    info = check_direct(harness_lib, oracle, x, y, z, L, 8.0)
    assert info[1] == 1                      # lanes over atoms
    assert info[0] < 2.2 * len(x) ** 2       # ~2 images per pair instead of the reference's 125
    check_direct(harness_lib, oracle, x, y, z, L, 10.0, cap=20)
    check_direct(harness_lib, oracle, x, y, z, L, 5.0, cap=12)
    for nm, (x, y, z, L, _, _) in c1_structures.items():
        check_direct(harness_lib, oracle, x, y, z, L, 10.0)
        check_direct(harness_lib, oracle, x, y, z, L, 8.0, cap=20)


def test_direct_windows_triclinic_adversarial_and_exact_cutoff(harness_lib, oracle, syn):
    b = syn.c3_batch(60)
    modes = set()
    for s in range(b.n_structs):
        info = check_direct(harness_lib, oracle, *b.structure(s), 8.0, algo=oracle.ALGO_BRUTE)
        modes.add(int(info[1]))
    assert 2 in modes                        # tiny cells take the lanes-over-images traversal
    adv = syn.adversarial_cells()
    for s in range(adv.n_structs):
        check_direct(harness_lib, oracle, *adv.structure(s), 8.0, algo=oracle.ALGO_BRUTE)
        check_direct(harness_lib, oracle, *adv.structure(s), 3.0, cap=5, algo=oracle.ALGO_BRUTE)
    sc = syn.simple_cubic(2.0, 3)            # ideal lattice: neighbours at exactly r_c, decided by rounding
    for rc in (4.0, 2.0, 2.0 * np.sqrt(2.0), 6.5):
        check_direct(harness_lib, oracle, *sc.structure(0), rc, algo=oracle.ALGO_BRUTE)
    check_direct(harness_lib, oracle, *sc.structure(0), 2.5, eps=0.0, algo=oracle.ALGO_BRUTE)
    one = syn.simple_cubic(1.0, 1)           # single atom, pure self-image edges
    check_direct(harness_lib, oracle, *one.structure(0), 3.05, algo=oracle.ALGO_BRUTE)


def test_direct_windows_unwrapped_positions(harness_lib, oracle, syn):
    """Rule R2: inputs are used as given.  Shift atoms by whole lattice vectors: windows follow, image offsets are
    relative to the stored positions."""
    x, y, z, L, _ = syn.perovskite_supercell(3, 3, 3, seed=5)
    rng = np.random.default_rng(5)
    k = rng.integers(-2, 3, size=(len(x), 3)).astype(np.float64)
    sh = k @ L
    check_direct(harness_lib, oracle, x + sh[:, 0], y + sh[:, 1], z + sh[:, 2], L, 6.0, algo=oracle.ALGO_BRUTE)
