"""GPU parity tests (run on the B200 box: pytest -m gpu).  Everything goes through the C ABI (libdgnn_b200.so).

Bar (BASELINE.json north_star): edge sets (src, dst, image) and distances BIT-EXACT vs the oracle after the canonical
row sort; near-cutoff pairs reported separately and equal; RBF features within 1e-5 relative (fp32 vs fp64 oracle).
"""
import hashlib
import json
import os

import numpy as np
import pytest

from conftest import assert_graph_equal

pytestmark = pytest.mark.gpu

RBF_RTOL = 1e-5  # north_star tolerance for the fp32 features


def build_and_fetch(gb, b, rc, cap, rbf_rc=None, dr=None, eps=1e-10, device_inputs=False):
    g = gb.build(b.atom_offsets, b.x, b.y, b.z, b.lattices, r_cutoff=rc, max_neighbors=cap, epsilon=eps,
                 rbf_r_cutoff=rbf_rc if rbf_rc else rc, dr=dr, skip_rbf=dr is None, device_inputs=device_inputs)
    return gb.to_host(g)


def check_batch(gb, oracle, b, rc, cap, dr=None, eps=1e-10, algo=None, rbf_rc=None):
    cap_o = oracle.UNCAPPED if cap is None else cap
    h = build_and_fetch(gb, b, rc, cap_o, rbf_rc, dr, eps)
    e0 = 0
    near = ties = 0
    maxdeg = 0
    for s in range(b.n_structs):
        x, y, z, L = b.structure(s)
        o = oracle.neighbor_list(x, y, z, L, rc, cap_o, eps, oracle.ALGO_CELLLIST if algo is None else algo)
        a0, a1 = int(b.atom_offsets[s]), int(b.atom_offsets[s + 1])
        assert h.struct_edge_offsets[s] == e0
        e1 = assert_graph_equal(h, e0, a0, a1, o, what=f"structure {s}")
        if dr is not None and o.n_edges:
            ref = oracle.edge_attr(o.dist, rbf_rc if rbf_rc else rc, dr)
            rel = np.max(np.abs(h.rbf[e0:e1].astype(np.float64) - ref) / ref)
            assert rel <= RBF_RTOL, f"structure {s}: RBF rel err {rel:.3e}"
        near += o.n_near_cutoff
        ties += o.n_tie_rows
        e0 = e1
    assert e0 == h.n_edges == h.struct_edge_offsets[-1]
    assert h.n_near_cutoff == near
    assert h.n_tie_rows == ties
    return h


def test_c2_perovskite_batch_uncapped_with_rbf(gpu_builder, oracle, syn):
    b = syn.c2_batch(24)
    h = check_batch(gpu_builder, oracle, b, 8.0, None, dr=0.2)
    assert h.K == 40 and h.n_edges > 24 * 25_000
    info = gpu_builder.info()
    assert info.n_cell_structs == 0 and info.kernel_launches >= 4     # supercells take the direct path
    assert info.onepass == 1                                           # small batch: single-pass build


def test_c2_perovskite_batch_cell_list_path(gpu_builder_cell, oracle, syn):
    b = syn.c2_batch(12, seed0=4321)
    h = check_batch(gpu_builder_cell, oracle, b, 8.0, None, dr=0.2)
    info = gpu_builder_cell.info()
    assert info.n_cell_structs == 12 and info.kernel_launches >= 10
    check_batch(gpu_builder_cell, oracle, b, 10.0, 20, dr=0.1)


def test_mixed_direct_and_cell_list_batch(pkg, oracle, syn):
    """One batch, both paths: structures above direct_max_atoms go through the cell list, the rest direct; the CSR
    is one array ordered by atom."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    none = (np.zeros(0), np.zeros(0), np.zeros(0), np.diag([5.0, 5.0, 5.0]))
    parts = [syn.perovskite_supercell(4, 4, 4, seed=4)[:4], syn.triclinic_cell(7)[:4], none,
             syn.perovskite_supercell(4, 4, 2, seed=9)[:4], syn.perovskite_supercell(5, 4, 3, seed=10)[:4],
             syn.triclinic_cell(8)[:4]]
    b = syn.concat(parts)
    gb = pkg.GraphBuilder(0)
    gb.set_option("direct_max_atoms", 200)
    check_batch(gb, oracle, b, 8.0, None, dr=0.2)
    assert gb.info().n_cell_structs == 2
    check_batch(gb, oracle, b, 6.0, 12, dr=0.2)
    gb.close()


def test_c2_capped_and_viewer_params(gpu_builder, oracle, syn):
    b = syn.c2_batch(8, seed0=5000)
    h = check_batch(gpu_builder, oracle, b, 10.0, 20, dr=0.1)       # reference defaults, K = 100
    assert h.K == 100 and h.n_edges == 20 * b.n_atoms
    check_batch(gpu_builder, oracle, b, 5.0, 12)                     # viewer: buildGraph(5.0, 12), no RBF


def test_c1_shipped_poscars_all_param_sets(gpu_builder, oracle, syn, c1_structures, golden_dir):
    exp = json.load(open(os.path.join(golden_dir, "c1_expected.json")))
    names = sorted(c1_structures)
    b = syn.concat([c1_structures[n][:4] for n in names])
    for pname, rc, cap, dr in (("reference_default", 10.0, 20, 0.1), ("viewer_default", 5.0, 12, None),
                               ("benchmark_uncapped", 8.0, None, 0.2), ("topology_uncapped", 10.0, None, 0.1)):
        h = check_batch(gpu_builder, oracle, b, rc, cap, dr=dr)
        # and against the committed golden digests (independent of the oracle binary on this box)
        for s, nm in enumerate(names):
            a0, a1 = int(b.atom_offsets[s]), int(b.atom_offsets[s + 1])
            e0, e1 = int(h.struct_edge_offsets[s]), int(h.struct_edge_offsets[s + 1])
            assert e1 - e0 == exp[nm][pname]["n_edges"]
            hh = hashlib.sha256()
            hh.update((h.row_ptr[a0:a1 + 1] - h.row_ptr[a0]).astype(np.int64).tobytes())
            hh.update(h.dst[e0:e1].tobytes())
            hh.update(np.ascontiguousarray(h.image[e0:e1]).tobytes())
            hh.update(h.dist[e0:e1].tobytes())
            assert hh.hexdigest() == exp[nm][pname]["edge_sha256"], (nm, pname)


def test_c3_triclinic_short_vectors(gpu_builder, gpu_builder_cell, oracle, syn):
    b = syn.c3_batch(300)
    check_batch(gpu_builder_cell, oracle, syn.c3_batch(60, seed0=99), 8.0, None, dr=0.2)
    h = check_batch(gpu_builder, oracle, b, 8.0, None, dr=0.2)
    # multi-image and self-image edges are present (same (src,dst) through several images; dst == src)
    src = np.repeat(np.arange(b.n_atoms), np.diff(h.row_ptr))
    local_src = src - np.repeat(b.atom_offsets[:-1], np.diff(h.struct_edge_offsets))
    assert np.any(local_src == h.dst)


def test_concurrent_contexts_on_one_device(pkg, oracle, syn):
    """Several contexts driven by their own host threads on ONE device (graph::ShardedBuilder's shape, the streamed
    build's twin): batches of different shapes -- different shared-memory sizes of the same kernels -- built side by side
    must neither fail nor differ from the same builds done alone.  (The per-function opt-in shared-memory attribute is
    set to the device maximum: a per-launch value raced between threads and a launch could fail with "invalid
    argument".)"""
    import threading
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    cases = [(syn.c2_batch(6, seed0=21), 8.0, None), (syn.c3_batch(40, seed0=5), 8.0, None),
             (syn.c2_batch(3, seed0=9), 10.0, 20), (syn.c3_batch(25, seed0=77), 6.0, 12)]
    ref = []
    gb0 = pkg.GraphBuilder(0)
    for b, rc, cap in cases:
        h = build_and_fetch(gb0, b, rc, oracle.UNCAPPED if cap is None else cap, None, 0.2, 1e-10)
        ref.append((h.n_edges, h.dst.copy(), h.dist.copy()))
    gb0.close()
    errors = []

    def worker(k):
        try:
            gb = pkg.GraphBuilder(0)
            b, rc, cap = cases[k]
            for _ in range(25):
                h = build_and_fetch(gb, b, rc, oracle.UNCAPPED if cap is None else cap, None, 0.2, 1e-10)
                assert h.n_edges == ref[k][0]
                assert np.array_equal(h.dst, ref[k][1]) and np.array_equal(h.dist, ref[k][2])
            gb.close()
        except Exception as e:  # noqa: BLE001
            errors.append((k, repr(e)))

    threads = [threading.Thread(target=worker, args=(k,)) for k in range(len(cases))]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors


def test_skewed_degrees_split_the_fill_pass(pkg, oracle, syn):
    """A few dense rows must not size every warp's row buffer: rows beyond the full-occupancy buffer are filled by a
    second launch (dgnn_build_info.split_big_rows), the rest by the first; both equal the oracle, and the single-launch
    plan (tune.split_rows = 0) gives the same graphs."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    c3 = syn.c3_batch(400, seed0=4242)
    mixed = syn.concat([syn.c2_batch(3, seed0=5).structure(s) for s in range(3)] +
                       [syn.triclinic_cell(9000 + s)[:4] for s in range(12)])
    for split in (1, 0):
        gb = pkg.GraphBuilder(0)
        gb.set_option("onepass", 0)
        gb.set_option("tune.split_rows", split)
        seen = []
        for b, rc, cap in ((c3, 8.0, None), (mixed, 8.0, None), (c3, 8.0, 400), (c3, 6.0, None)):
            h = check_batch(gb, oracle, b, rc, cap, dr=0.2)
            seen.append((gb.info().split_big_rows, gb.info().row_cap, int(np.diff(h.row_ptr).max())))
        # the same skewed cells next to structures on the cell-list path (whose rows keep the buffer for the largest row)
        gb.set_option("direct_max_atoms", 100)
        both = syn.concat([c3.structure(s) for s in range(150)] + [syn.c2_batch(2, seed0=8).structure(s) for s in range(2)])
        check_batch(gb, oracle, both, 8.0, None, dr=0.2)
        assert gb.info().n_cell_structs == 2
        if split:
            assert gb.info().split_big_rows > 0
            assert seen[0][0] > 0 and seen[0][1] < seen[0][2]      # C3: max degree ~800, buffer of the first launch ~350
            assert all(n * 8 <= c3.n_atoms for n, _, _ in seen)
        else:
            assert all(n == 0 for n, _, _ in seen)
        gb.close()


def test_adversarial_cells_reference_image_range(gpu_builder, gpu_builder_cell, oracle, syn):
    b = syn.adversarial_cells()
    for gb in (gpu_builder, gpu_builder_cell):
        check_batch(gb, oracle, b, 8.0, None, dr=0.2, algo=oracle.ALGO_BRUTE)
        check_batch(gb, oracle, b, 3.0, 5, algo=oracle.ALGO_BRUTE)


@pytest.mark.parametrize("path", ["direct", "cell"])
def test_known_answers_and_edge_cases(gpu_builder, gpu_builder_cell, oracle, syn, pkg, path):
    gb = gpu_builder if path == "direct" else gpu_builder_cell
    sc = syn.simple_cubic(1.0, 1)
    h = check_batch(gb, oracle, sc, 3.05, None, algo=oracle.ALGO_BRUTE)
    d2 = np.round(h.dist ** 2).astype(int)
    assert {k: int((d2 == k).sum()) for k in (1, 2, 3, 4, 5, 6, 8, 9)} == {1: 6, 2: 12, 3: 8, 4: 6, 5: 24, 6: 24, 8: 12, 9: 30}
    # neighbours at exactly r_c: strict '<' decided by rounding, identical to the oracle
    sc3 = syn.simple_cubic(2.0, 3)
    check_batch(gb, oracle, sc3, 4.0, None, algo=oracle.ALGO_BRUTE)
    check_batch(gb, oracle, sc3, 2.0, None, algo=oracle.ALGO_BRUTE)
    # epsilon semantics (R6)
    check_batch(gb, oracle, sc3, 2.5, None, eps=0.0, algo=oracle.ALGO_BRUTE)
    check_batch(gb, oracle, sc3, 6.5, None, eps=6.1, algo=oracle.ALGO_BRUTE)
    # empty batch, empty structures, ragged sizes
    empty = syn.concat([])
    g = gb.build(empty.atom_offsets, empty.x, empty.y, empty.z, empty.lattices, 8.0, 20, dr=0.2, rbf_r_cutoff=8.0)
    assert g.n_edges == 0 and g.n_structs == 0
    one = syn.perovskite_supercell(2, 2, 2, seed=3)[:4]
    none = (np.zeros(0), np.zeros(0), np.zeros(0), np.diag([5.0, 5.0, 5.0]))
    tiny = syn.triclinic_cell(7)[:4]
    ragged = syn.concat([none, one, none, tiny, syn.perovskite_supercell(4, 4, 4, seed=4)[:4], none])
    h = check_batch(gb, oracle, ragged, 8.0, None, dr=0.2)
    assert h.struct_edge_offsets[0] == h.struct_edge_offsets[1] == 0
    # max_neighbors = 0 and 1
    check_batch(gb, oracle, ragged, 8.0, 1)
    h0 = build_and_fetch(gb, ragged, 8.0, 0)
    assert h0.n_edges == 0


def test_error_codes(gpu_builder, syn, pkg):
    gb = gpu_builder
    x, y, z, L = syn.perovskite_supercell(2, 2, 2, seed=3)[:4]
    b = syn.concat([(x, y, z, L)])
    with pytest.raises(pkg.DgnnError) as ei:
        gb.build(b.atom_offsets, b.x, b.y, b.z, b.lattices, r_cutoff=-1.0)
    assert ei.value.code == -1
    bad = syn.concat([(x, y, z, np.array([[1.0, 0, 0], [2.0, 0, 0], [0, 0, 1.0]]))])
    with pytest.raises(pkg.DgnnError) as ei:
        gb.build(bad.atom_offsets, bad.x, bad.y, bad.z, bad.lattices, r_cutoff=8.0)
    assert ei.value.code == -4  # singular lattice
    thin = syn.concat([(x, y, z, np.diag([0.05, 8.0, 8.0]))])
    with pytest.raises(pkg.DgnnError) as ei:
        gb.build(thin.atom_offsets, thin.x, thin.y, thin.z, thin.lattices, r_cutoff=8.0)
    assert ei.value.code in (-5, -6)
    # the context stays usable after an error
    g = gb.build(b.atom_offsets, b.x, b.y, b.z, b.lattices, r_cutoff=5.0, max_neighbors=12, skip_rbf=True)
    assert g.n_edges == 12 * len(x)


def test_block_decomposition_and_unstaged_paths(pkg, oracle, syn):
    """Medium supercell: (i) default plan (several work items), (ii) tiny tiles, (iii) forced unstaged traversal --
    all three must give the same bit-exact result."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    x, y, z, L, _ = syn.perovskite_supercell(9, 8, 7, seed=11)
    b = syn.concat([(x, y, z, L)])
    o = oracle.neighbor_list(x, y, z, L, 8.0, oracle.UNCAPPED, 1e-10, oracle.ALGO_CELLLIST)
    # default = fp32 tiles staged in shared memory (dgnn_cell32.cu); cell32 = 0 = the fp64 traversal (staged with
    # tile_limit > 0, else straight from the L2-resident cloud), with and without hit lists
    for opts in ({}, {"tune.tile_target": 512, "tune.src_target": 48}, {"tune.tile_target": 4096, "tune.src_target": 1024}, {"cell_frac": 0.3},
                 {"cell_frac": 1.0}, {"cell_frac": 0.3, "tune.tile_target": 256, "tune.src_target": 16},
                 {"cell32": 0}, {"cell32": 0, "tune.tile_limit": 4096}, {"cell32": 0, "tune.tile_limit": 4096, "tune.tile_target": 512, "tune.src_target": 48},
                 {"cell32": 0, "tune.tile_limit": 4096, "hit_cap": 0}, {"cell32": 0, "hit_cap": 0}, {"cell32": 0, "hit_cap": 160},
                 {"cell32": 0, "tune.tile_limit": 4096, "hit_cap": 160}, {"cell32": 0, "cell_frac": 0.3},
                 {"cell32": 0, "cell_frac": 1.0}, {"direct_max_atoms": 4096}):
        gb = pkg.GraphBuilder(0)
        if "direct_max_atoms" not in opts:
            gb.set_option("direct_max_atoms", 0)
        for k, v in opts.items():
            gb.set_option(k, v)
        g = gb.build(b.atom_offsets, b.x, b.y, b.z, b.lattices, 8.0, oracle.UNCAPPED, rbf_r_cutoff=8.0, dr=0.2)
        h = gb.to_host(g)
        assert_graph_equal(h, 0, 0, len(x), o, what=str(opts))
        info = gb.info()
        if "direct_max_atoms" not in opts:
            if opts.get("cell32", 1) == 0 and opts.get("tune.tile_limit", 0) == 0:
                assert info.staged_tile_cap == 0
            else:
                assert info.staged_tile_cap > 0
        if opts.get("tune.tile_target", 4096) <= 512:
            assert info.n_items > 20
        assert info.n_cell_structs == (0 if "direct_max_atoms" in opts else 1)
        gb.close()


def test_device_resident_inputs_and_rbf_standalone(gpu_builder, oracle, syn):
    import torch
    gb = gpu_builder
    b = syn.c2_batch(5, seed0=777)
    dev = torch.device("cuda:0")
    t = [torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in (b.atom_offsets, b.x, b.y, b.z, b.lattices)]
    torch.cuda.synchronize()
    g = gb.build(*t, r_cutoff=8.0, max_neighbors=oracle.UNCAPPED, rbf_r_cutoff=8.0, dr=0.2, device_inputs=True)
    h_dev = gb.to_host(g)
    h_host = build_and_fetch(gb, b, 8.0, oracle.UNCAPPED, 8.0, 0.2)
    for f in ("row_ptr", "dst", "image", "dist", "rbf"):
        assert np.array_equal(getattr(h_dev, f), getattr(h_host, f)), f
    # stand-alone gaussian_rbf replacement, host and device pointers, fp32 row-major and fp64 column-major
    d = np.concatenate([[0.0, 4.0, 8.0 - 1e-9], np.random.default_rng(0).uniform(0.0, 8.0, 5000)])
    # K = 40, 100 (float4 rows); 26, 50, 11, 7, 1 (rows not 16-byte aligned: float4s straddle edges, partial tail)
    for rc, dr in ((8.0, 0.2), (10.0, 0.1), (8.0, 0.3), (10.0, 0.2), (8.0, 0.7), (7.0, 1.0), (8.0, 5.0)):
        ref = oracle.edge_attr(d, rc, dr)
        out = gb.rbf_expand(d, rc, dr)
        assert out.shape == ref.shape
        assert np.max(np.abs(out - ref) / ref) <= RBF_RTOL
        out64 = gb.rbf_expand(d, rc, dr, f64_colmajor=True)
        assert np.max(np.abs(out64 - ref) / ref) <= 1e-12
    dd = torch.from_numpy(d).to(dev)
    od = torch.empty((d.shape[0], 40), dtype=torch.float32, device=dev)
    gb.rbf_expand(dd, 8.0, 0.2, out=od)
    assert np.max(np.abs(od.cpu().numpy() - oracle.edge_attr(d, 8.0, 0.2)) / oracle.edge_attr(d, 8.0, 0.2)) <= RBF_RTOL


def test_determinism_and_symmetry_property(gpu_builder, oracle, syn):
    """Size-independent properties on a larger batch: two builds are byte-identical; every edge (i -> j, n) has its
    mirror (j -> i, -n) with the bit-identical distance; rows are sorted."""
    b = syn.c2_batch(200, seed0=9000)
    h1 = build_and_fetch(gpu_builder, b, 8.0, oracle.UNCAPPED, 8.0, 0.2)
    h2 = build_and_fetch(gpu_builder, b, 8.0, oracle.UNCAPPED, 8.0, 0.2)
    for f in ("row_ptr", "dst", "image", "dist", "rbf"):
        assert np.array_equal(getattr(h1, f), getattr(h2, f)), f
    counts = oracle.batch_edge_counts(b.atom_offsets, b.x, b.y, b.z, b.lattices, 8.0)
    assert np.array_equal(np.diff(h1.struct_edge_offsets), counts)
    src = np.repeat(np.arange(b.n_atoms, dtype=np.int64), np.diff(h1.row_ptr))
    sid = np.repeat(np.arange(b.n_structs), np.diff(h1.struct_edge_offsets))
    dstg = h1.dst.astype(np.int64) + b.atom_offsets[sid]
    n = h1.image[:, :3].astype(np.int64)
    key_f = ((src * b.n_atoms + dstg) * 8 + (n[:, 0] + 3)) * 64 + (n[:, 1] + 3) * 8 + (n[:, 2] + 3)
    key_b = ((dstg * b.n_atoms + src) * 8 + (3 - n[:, 0])) * 64 + (3 - n[:, 1]) * 8 + (3 - n[:, 2])
    # The strict '<' is decided on a rounded d2, and the mirrored pair is computed from (pos_j + off(-n)) - pos_i,
    # which rounds differently: pairs inside the 1e-6 A near-cutoff band (here: self-images at exactly c = 8 A) may
    # legitimately exist in one direction only -- in the reference too.  They are excluded from the symmetry check.
    far = np.abs(h1.dist - 8.0) > 1e-6
    assert h1.n_near_cutoff >= int((~far).sum())
    kf, kb, df = key_f[far], key_b[far], h1.dist[far]
    of, ob = np.argsort(kf), np.argsort(kb)
    assert np.array_equal(kf[of], kb[ob])
    assert np.allclose(df[of], df[ob], rtol=0, atol=1e-12)
    seg_start = np.zeros(h1.n_edges, dtype=bool)
    seg_start[h1.row_ptr[:-1][np.diff(h1.row_ptr) > 0]] = True
    dd = np.diff(h1.dist)
    assert np.all((dd >= 0) | seg_start[1:])


@pytest.mark.parametrize("hit_cap", [-1, 0, 64, 160, 4096])
def test_hit_list_and_traversal_fill_paths_agree(pkg, oracle, syn, hit_cap):
    """The fill pass has two producers: the count pass's hit lists (k_fill_hits) and the repeated traversal
    (k_direct<FILL>); rows that outgrow hit_cap take the latter.  Every split must give the same bit-exact graphs."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    gb = pkg.GraphBuilder(0)
    gb.set_option("hit_cap", hit_cap)
    b = syn.concat([syn.perovskite_supercell(4, 4, 2, seed=21)[:4], syn.triclinic_cell(3)[:4],
                    syn.perovskite_supercell(3, 3, 3, seed=22)[:4], syn.triclinic_cell(5)[:4],
                    syn.perovskite_supercell(4, 4, 4, seed=23)[:4]])
    check_batch(gb, oracle, b, 8.0, None, dr=0.2)
    info = gb.info()
    assert info.hit_cap == (min(hit_cap, 512) if hit_cap >= 0 else info.hit_cap)
    if hit_cap == 0:
        assert info.hit_cap == 0
    check_batch(gb, oracle, b, 10.0, 20, dr=0.1)
    check_batch(gb, oracle, b, 5.0, 12)
    gb.close()


@pytest.mark.parametrize("opts,expect", [
    ({"onepass": 0, "hit_cap": 0}, 0),                       # count + traversal fill
    ({"onepass": 0}, 0),                                      # count (+ hit lists where they pay) + fill
    ({"onepass": 1}, 1),                                      # single pass, row buffer / arena from the density estimate
    ({"onepass": 1, "tune.onepass_row_cap": 64}, 1),               # most rows outgrow the row buffer -> completed by a fill pass
    ({"onepass": 1, "tune.onepass_arena_scale": 0.25}, 2),         # arena estimate too small -> every row rewritten
    ({"onepass": 1, "tune.lb_spin_limit": 1}, 3),                  # look-back gives up at once -> rebuilt with count + fill
])
def test_single_pass_build_matches_two_pass(pkg, oracle, syn, opts, expect):
    """The single-pass build (k_direct<ONE>: look-back row offsets, no count pass) and its two recovery routes must
    give the same bit-exact CSR, statistics and features as count + fill."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    gb = pkg.GraphBuilder(0)
    for k, v in opts.items():
        gb.set_option(k, v)
    none = (np.zeros(0), np.zeros(0), np.zeros(0), np.diag([5.0, 5.0, 5.0]))
    b = syn.concat([syn.perovskite_supercell(4, 4, 2, seed=31)[:4], none, syn.triclinic_cell(13)[:4],
                    syn.perovskite_supercell(3, 3, 3, seed=32)[:4], syn.triclinic_cell(15)[:4], none,
                    syn.perovskite_supercell(4, 4, 4, seed=33)[:4], syn.perovskite_supercell(2, 2, 2, seed=34)[:4]])
    check_batch(gb, oracle, b, 8.0, None, dr=0.2)
    info = gb.info()
    # (a look-back only times out if it has to wait at all: with the spin limit at 1 that is near certain on the many
    # chunks of the larger batch below, not guaranteed on this small one)
    assert info.onepass == expect or (expect == 3 and info.onepass == 1)
    if opts.get("tune.onepass_row_cap"):
        assert info.onepass_fallback_rows > 0
    check_batch(gb, oracle, b, 10.0, 20, dr=0.1)           # capped rows, K = 100
    check_batch(gb, oracle, b, 8.0, None, dr=0.3)           # K = 26: separate RBF launch (no float4 epilogue)
    check_batch(gb, oracle, b, 5.0, 12)                     # no features
    check_batch(gb, oracle, b, 8.0, 1)
    h0 = build_and_fetch(gb, b, 8.0, 0)
    assert h0.n_edges == 0
    # a larger homogeneous batch: many chunks per CTA, look-back across CTAs
    big = syn.c2_batch(96, seed0=31000)
    seen = set()
    for _ in range(3):
        check_batch(gb, oracle, big, 8.0, None, dr=0.2)
        seen.add(gb.info().onepass)
    if expect == 3:
        assert 3 in seen, "the timed-out look-back must have been redone with count + fill at least once"
    gb.close()


@pytest.mark.parametrize("opts", [{"upload_chunks": 3}, {"upload_chunks": 8, "onepass": 0}, {"upload_chunks": 5, "onepass": 0, "hit_cap": 0},
                                  {"upload_chunks": 2, "direct_max_atoms": 150}])
def test_chunked_upload_overlaps_count_pass(pkg, oracle, syn, opts):
    """Host inputs travel in structure-aligned chunks on a copy stream while the prep + count pass of earlier chunks
    run (dgnn_build, host-input path); chunk boundaries fall on empty structures, tiny and large ones."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    gb = pkg.GraphBuilder(0)
    gb.set_option("tune.upload_min_atoms", 0)
    for k, v in opts.items():
        gb.set_option(k, v)
    none = (np.zeros(0), np.zeros(0), np.zeros(0), np.diag([5.0, 5.0, 5.0]))
    parts = [none, syn.triclinic_cell(31)[:4], syn.perovskite_supercell(4, 4, 2, seed=51)[:4], none, none,
             syn.perovskite_supercell(3, 3, 3, seed=52)[:4], syn.triclinic_cell(33)[:4],
             syn.perovskite_supercell(2, 2, 2, seed=53)[:4], syn.perovskite_supercell(4, 4, 2, seed=54)[:4], none]
    b = syn.concat(parts)
    check_batch(gb, oracle, b, 8.0, None, dr=0.2)
    check_batch(gb, oracle, b, 10.0, 20, dr=0.1)
    # pinned host tensors (the bench's e2e leg) and plain numpy give the same graphs
    t = [torch.from_numpy(np.ascontiguousarray(a)).pin_memory() for a in (b.atom_offsets, b.x, b.y, b.z, b.lattices)]
    g = gb.build(*t, r_cutoff=8.0, max_neighbors=oracle.UNCAPPED, rbf_r_cutoff=8.0, dr=0.2)
    h1 = gb.to_host(g)
    h2 = build_and_fetch(gb, b, 8.0, oracle.UNCAPPED, 8.0, 0.2)
    for f in ("row_ptr", "dst", "image", "dist", "rbf"):
        assert np.array_equal(getattr(h1, f), getattr(h2, f)), f
    gb.close()


@pytest.mark.parametrize("onepass", [0, 1])
def test_offset_sinks_receive_row_ptr_during_the_build(pkg, oracle, syn, onepass):
    """dgnn_set_offset_sinks: row_ptr / struct_edge_offsets are in the host sinks when dgnn_build returns (copied on the
    copy stream while the fill runs for the two-pass plans, at the end for the single-pass plan), for host and device
    inputs, direct and cell-list structures, and stop being written once the sinks are cleared."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    gb = pkg.GraphBuilder(0)
    gb.set_option("onepass", onepass)
    gb.set_option("tune.upload_min_atoms", 0)
    none = (np.zeros(0), np.zeros(0), np.zeros(0), np.diag([5.0, 5.0, 5.0]))
    parts = [syn.perovskite_supercell(4, 4, 2, seed=61)[:4], none, syn.triclinic_cell(62)[:4],
             syn.perovskite_supercell(3, 3, 3, seed=63)[:4], syn.perovskite_supercell(2, 2, 2, seed=64)[:4]]
    b = syn.concat(parts)
    S, N = len(parts), b.n_atoms
    for rc, cap in ((8.0, oracle.UNCAPPED), (10.0, 20)):
        ref = build_and_fetch(gb, b, rc, cap, rc, 0.2)
        h_rp = torch.full((N + 1,), -1, dtype=torch.int64).pin_memory()
        h_so = torch.full((S + 1,), -1, dtype=torch.int64).pin_memory()
        gb.set_offset_sinks(h_rp, h_so)
        for dev_in in (False, True):
            h_rp.fill_(-1), h_so.fill_(-1)
            arrs = [np.ascontiguousarray(a) for a in (b.atom_offsets, b.x, b.y, b.z, b.lattices)]
            if dev_in:
                arrs = [torch.from_numpy(a).cuda() for a in arrs]
            g = gb.build(*arrs, r_cutoff=rc, max_neighbors=cap, rbf_r_cutoff=rc, dr=0.2, device_inputs=dev_in)
            assert np.array_equal(h_rp.numpy(), ref.row_ptr), (rc, cap, dev_in)   # no sync, no copy_to_host needed
            assert np.array_equal(h_so.numpy(), ref.struct_edge_offsets)
            assert int(h_rp[-1]) == int(g.n_edges)
        gb.set_offset_sinks(None, None)
        h_rp.fill_(-7)
        build_and_fetch(gb, b, rc, cap, rc, 0.2)
        assert int(h_rp[0]) == -7 and int(h_rp[-1]) == -7
    gb.close()


@pytest.mark.parametrize("opts", [{"onepass": 0}, {"onepass": 1}, {"onepass": 0, "hit_cap": 0}],
                         ids=["hit-lists", "single-pass", "traversal-fill"])
def test_truncated_rows_topk_prefilter(pkg, oracle, syn, opts):
    """Capped rows are not sorted in full: a d2 threshold keeps a little more than max_neighbors + 1 candidates, which
    are compacted and sorted.  Caps from 1 to most of the row, jittered and ideal lattices (whole shells of equal
    distances at the threshold and at the truncation), skewed cells -- rows, tie order and the tie statistic must equal
    the exhaustive oracle in every build plan."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    gb = pkg.GraphBuilder(0)
    for k, v in opts.items():
        gb.set_option(k, v)
    batches = [syn.concat([syn.perovskite_supercell(4, 4, 2, seed=1)[:4], syn.triclinic_cell(5)[:4]]),
               syn.concat([syn.perovskite_supercell(3, 3, 2, seed=9, jitter=0.0)[:4], syn.simple_cubic(2.0, 4).structure(0)])]
    for b in batches:
        for rc, dr in ((8.0, 0.2), (10.0, 0.25)):
            for cap in (1, 7, 20, 33, 64, 100):
                check_batch(gb, oracle, b, rc, cap, dr=dr, algo=oracle.ALGO_BRUTE)
    check_batch(gb, oracle, syn.adversarial_cells(), 8.0, 20, dr=0.2, algo=oracle.ALGO_BRUTE)
    gb.close()


@pytest.mark.parametrize("sym", [0, 1])
def test_symmetric_count_pass_is_exact(pkg, oracle, syn, sym):
    """The symmetric count pass (one sweep decides i -> j and j -> i, the near-cutoff band recomputes the mirror) must
    give the degrees and the near-cutoff statistic of the two-sided sweep: lattice sites exactly at r_c, skewed cells,
    tiny cells with many images, coincident atoms, far-from-origin (unwrapped) coordinates."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    gb = pkg.GraphBuilder(0)
    for k, v in (("onepass", 0), ("hit_cap", 0), ("sym_count", sym)):
        gb.set_option(k, v)
    sc3 = syn.simple_cubic(2.0, 3)
    for rc in (4.0, 2.0, 2.0 * np.sqrt(2.0), 6.0):           # shells exactly at the cutoff: strict '<' on rounded d2
        check_batch(gb, oracle, sc3, rc, None, algo=oracle.ALGO_BRUTE)
    check_batch(gb, oracle, syn.adversarial_cells(), 8.0, None, dr=0.2, algo=oracle.ALGO_BRUTE)
    check_batch(gb, oracle, syn.c3_batch(80, seed0=7), 8.0, None, dr=0.2)
    check_batch(gb, oracle, syn.c2_batch(6, seed0=123), 8.0, None, dr=0.2)
    check_batch(gb, oracle, syn.c2_batch(4, seed0=321), 10.0, 20, dr=0.1)
    # coincident atoms and an unwrapped, far-away copy of a structure (positions are never wrapped, rule R2)
    x, y, z, L = syn.perovskite_supercell(2, 2, 2, seed=3)[:4]
    x2, y2, z2 = np.append(x, x[5]), np.append(y, y[5]), np.append(z, z[5])
    shift = 7.0 * L[0] - 5.0 * L[1] + 1000.0 * L[2]
    far = (x + shift[0], y + shift[1], z + shift[2], L)
    check_batch(gb, oracle, syn.concat([(x2, y2, z2, L), far]), 6.0, None, dr=0.2, algo=oracle.ALGO_BRUTE)
    # lattice shells EXACTLY at r_c with unwrapped coordinates around 1e3 A (the mirror-band logic of the symmetric count
    # and any reduced-precision pre-filter must leave these to the fp64 reference arithmetic)
    b3 = syn.simple_cubic(2.0, 3)
    for sh in ((1024.0, -2048.0, 512.0), (1000.0, 3000.0, -1500.0)):
        moved = (b3.x + sh[0], b3.y + sh[1], b3.z + sh[2], b3.lattices[0].reshape(3, 3))
        for rc in (4.0, 2.0 * np.sqrt(2.0), 6.0, np.nextafter(6.0, 7.0)):
            check_batch(gb, oracle, syn.concat([moved]), rc, None, algo=oracle.ALGO_BRUTE)
    rng = np.random.default_rng(3)
    x, y, z, L = syn.perovskite_supercell(3, 3, 2, seed=8)[:4]
    for _ in range(3):
        n = rng.integers(-60, 60, size=3).astype(np.float64)
        sh = n @ L
        check_batch(gb, oracle, syn.concat([(x + sh[0], y + sh[1], z + sh[2], L)]), 8.0, None, dr=0.2)
    gb.close()




def test_tie_probe_selects_the_tie_ranking(pkg, oracle, syn):
    """The uncapped sweep-fill kernel has two tie rankings (compact per entry / warp-cooperative); sampled rows decide
    which one a batch gets (dgnn_tie_probe.cu).  Thermal structures: no ties, the compact one; ideal lattices (whole
    coordination shells on one sort key): the cooperative one.  Rows, tie order and the tie statistic equal the oracle
    either way and with the choice forced."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    ideal = syn.concat([syn.perovskite_supercell(4, 4, 2, seed=s, jitter=0.0)[:4] for s in range(6)])
    thermal = syn.c2_batch(6, seed0=77)
    for probe in (1, 0, 2):
        gb = pkg.GraphBuilder(0)
        for k, v in (("onepass", 0), ("hit_cap", 0), ("tune.tie_probe", probe)):
            gb.set_option(k, v)
        check_batch(gb, oracle, thermal, 8.0, None, dr=0.2)
        it = gb.info()
        check_batch(gb, oracle, ideal, 8.0, None, dr=0.2, algo=oracle.ALGO_BRUTE)
        ii = gb.info()
        if probe == 1:
            assert 0.99 <= it.tie_run_mean < 1.5 and it.full_tie_sort == 0
            assert ii.tie_run_mean >= 3.0 and ii.full_tie_sort == 1
        else:
            assert it.tie_run_mean == 0.0 and ii.tie_run_mean == 0.0          # not probed
            assert it.full_tie_sort == ii.full_tie_sort == (1 if probe == 2 else 0)
        # capped rows never take the lean kernel: the flag stays 0
        check_batch(gb, oracle, ideal, 8.0, 20, dr=0.2, algo=oracle.ALGO_BRUTE)
        assert gb.info().full_tie_sort == 0
        gb.close()


@pytest.mark.parametrize("onepass", [0, 1])
def test_long_tie_runs_cross_the_ranking_window(pkg, oracle, syn, onepass):
    """Coordination shells of more than 32 (and more than 64) exactly or nearly equal distances: the warp-cooperative
    ranking works on 32 sorted positions at a time and counts the rest of such a run block by block
    (dg_tie_block_count: groups of equal distance by size, the own group by a radix walk over dst).  Simple-cubic
    supercells (shells of up to 72+ sites, every site a different dst or another image of the same one) and ideal
    perovskite cells at r_c = 10 / 12, uncapped and truncated in the middle of a shell, against the exhaustive oracle."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")

    def cubic(n, a):
        g = np.arange(n) * a
        X, Y, Z = np.meshgrid(g, g, g, indexing="ij")
        return X.ravel().astype(np.float64), Y.ravel().astype(np.float64), Z.ravel().astype(np.float64), np.diag([n * a] * 3)

    def fcc(n, a):  # the shell at r^2 = 6.5 a^2 has 72 sites (320 sites within it)
        g = np.arange(n)
        C = np.stack(np.meshgrid(g, g, g, indexing="ij"), axis=-1).reshape(-1, 1, 3).astype(np.float64)
        B = np.array([[0, 0, 0], [0.5, 0.5, 0], [0.5, 0, 0.5], [0, 0.5, 0.5]], dtype=np.float64)[None]
        P = ((C + B) * a).reshape(-1, 3)
        return P[:, 0].copy(), P[:, 1].copy(), P[:, 2].copy(), np.diag([n * a] * 3)

    sc = syn.concat([cubic(4, 2.0), cubic(3, 1.7), cubic(5, 2.5)])
    fc = syn.concat([fcc(3, 3.0), fcc(2, 3.0)])
    ideal = syn.concat([syn.perovskite_supercell(4, 4, 2, seed=s, jitter=0.0)[:4] for s in range(2)])
    gb = pkg.GraphBuilder(0)
    for k, v in (("onepass", onepass), ("hit_cap", 0), ("tune.tie_probe", 2)):
        gb.set_option(k, v)
    longest = 0
    for b, rc, cap in ((sc, 6.0, None), (sc, 7.5, None), (sc, 6.0, 50), (sc, 7.5, 129), (fc, 7.8, None), (fc, 7.8, 280),
                       (fc, 7.8, 300), (ideal, 10.0, None), (ideal, 12.0, None), (ideal, 10.0, 100)):
        h = check_batch(gb, oracle, b, rc, cap, dr=0.25, algo=oracle.ALGO_BRUTE)
        # longest run of equal rounded distances in any row of this build (what the ranking had to order)
        for a in range(0, b.n_atoms, 37):
            d = np.round(h.dist[h.row_ptr[a]:h.row_ptr[a + 1]], 9)
            if d.size:
                longest = max(longest, int(np.unique(d, return_counts=True)[1].max()))
    assert longest > 64
    gb.close()


def test_thresholded_hit_lists_of_truncated_rows(pkg, oracle, syn):
    """Truncated rows (max_neighbors << degree): the count pass stores only accepted pairs below a per-structure d2
    threshold sized from the density for a little more than max_neighbors + 1 of them, and the fill pass sorts those.
    Rows the threshold cannot serve -- sparser than the structure's mean (atoms above a slab, next to a void), or
    denser than the slots -- must be completed by the traversal fill; whole shells of equal distances at the threshold
    (ideal lattices) must not lose a tie partner.  Every case equals the exhaustive oracle."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    x, y, z, L = syn.perovskite_supercell(4, 4, 2, seed=5)[:4]
    half = z < 0.5 * L[2, 2]
    lone = np.zeros_like(half)
    lone[np.flatnonzero(~half)[:3]] = True                      # three atoms above a slab: sparse rows
    keep = half | lone
    slab = (x[keep], y[keep], z[keep], L)
    tall = L.copy()
    tall[2] *= 3.0                                              # mostly vacuum: local density 3x the mean
    vac = (x, y, z, tall)
    thermal = syn.c2_batch(5, seed0=11)
    ideal = syn.concat([syn.perovskite_supercell(4, 4, 2, seed=s, jitter=0.0)[:4] for s in range(3)])
    mixed = syn.concat([slab, vac, syn.perovskite_supercell(3, 3, 2, seed=2)[:4], syn.triclinic_cell(5)[:4]])
    for thr in (1, 0):
        gb = pkg.GraphBuilder(0)
        gb.set_option("onepass", 0)
        gb.set_option("tune.hit_threshold", thr)
        caps_seen = []
        for b, algo in ((thermal, None), (ideal, oracle.ALGO_BRUTE), (mixed, oracle.ALGO_BRUTE)):
            for rc, cap in ((8.0, 20), (10.0, 20), (8.0, 1), (8.0, 12), (10.0, 40)):
                check_batch(gb, oracle, b, rc, cap, dr=0.2, algo=algo)
                caps_seen.append(gb.info().hit_cap)
        if thr:
            assert min(caps_seen) > 0 and max(caps_seen) <= 160      # slots follow the cap, not the degree
        else:
            assert max(caps_seen) >= 192
        gb.close()
