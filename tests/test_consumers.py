"""Consumers of the CSR (collation f1, viewer arrays f4, neighbourhood clouds f2).

CPU part: the numpy restatements in oracle/oracle_py.py are pinned on the C++ oracle's own displacement output and on
hand-checked small cases.  GPU part (-m gpu): the device kernels against those restatements, through the C ABI:
integer / index arrays bit-exact, float32 casts bit-exact (same fp64 value cast once), fp64 clouds bit-exact."""
import numpy as np
import pytest


def _emb(F=92, keys=range(0, 6), seed=7):
    rng = np.random.default_rng(seed)
    return {int(k): rng.standard_normal(F) for k in keys}


def test_displacement_restatement_matches_cpp_oracle(oracle, syn):
    """numpy (dst, image) -> displacement equals the displacement the exhaustive / KD-tree oracle computes while
    searching (cloud_pos - query_pos, neighbor_list.cpp:51), bit for bit."""
    for st in (syn.perovskite_supercell(2, 2, 2, seed=3)[:4], syn.triclinic_cell(5)[:4], syn.triclinic_cell(11)[:4]):
        x, y, z, L = st
        for algo in (oracle.ALGO_BRUTE, oracle.ALGO_KDTREE):
            o = oracle.neighbor_list(x, y, z, L, 6.0, oracle.UNCAPPED, 1e-10, algo)
            d = oracle.displacements(x, y, z, L, o)
            assert np.array_equal(d, o.disp)
            # and the distance is the norm of that displacement up to rounding
            assert np.allclose(np.sqrt((d ** 2).sum(1)), o.dist, rtol=1e-14, atol=0)


def test_collate_restatement_small_case(oracle, syn):
    x, y, z, L = syn.simple_cubic(2.0, 2).structure(0)
    o = oracle.neighbor_list(x, y, z, L, 2.5, oracle.UNCAPPED)
    emb = {0: np.arange(4.0), 1: np.arange(4.0) + 10}
    keys = np.array([0, 1] * (len(x) // 2))
    X, ei, batch = oracle.collate([o, o], [keys, keys], emb)
    n, e = len(x), o.n_edges
    assert X.shape == (2 * n, 4) and X.dtype == np.float32 and ei.dtype == np.int64
    assert np.array_equal(ei[:, :e], np.stack([np.repeat(np.arange(n), np.diff(o.row_ptr)), o.dst]))
    assert np.array_equal(ei[:, e:], ei[:, :e] + n)           # second graph: node offset
    assert np.array_equal(batch, np.repeat([0, 1], n))
    assert np.array_equal(X[1], emb[1].astype(np.float32))
    with pytest.raises(KeyError):                               # std::map::at -> out_of_range
        oracle.collate([o], [np.full(n, 7)], emb)


def test_viewer_and_cloud_restatements(oracle, syn):
    x, y, z, L = syn.perovskite_supercell(2, 2, 2, seed=5)[:4]
    o = oracle.neighbor_list(x, y, z, L, 5.0, 12)
    src, tgt, dist, disp = oracle.viewer_arrays(o)
    assert src.dtype == np.int32 and dist.dtype == np.float32 and disp.shape == (3 * o.n_edges,)
    assert np.array_equal(np.bincount(src, minlength=len(x)), np.diff(o.row_ptr))
    c = oracle.neighborhood_cloud(x, y, z, o, 3)
    assert c.shape == (13, 3) and np.array_equal(c[0], [x[3], y[3], z[3]])
    # every cloud point lies at the edge's distance from the centre
    assert np.allclose(np.linalg.norm(c[1:] - c[0], axis=1), o.dist[o.row_ptr[3]:o.row_ptr[4]], rtol=1e-12)


# ---------------------------------------------------------------------------------------------------------------------
def _mixed_batch(syn):
    none = (np.zeros(0), np.zeros(0), np.zeros(0), np.diag([5.0, 5.0, 5.0]))
    return syn.concat([syn.perovskite_supercell(3, 3, 2, seed=41)[:4], none, syn.triclinic_cell(21)[:4],
                       syn.perovskite_supercell(4, 4, 2, seed=42)[:4], syn.triclinic_cell(23)[:4]])


def _oracle_graphs(oracle, b, rc, cap):
    return [oracle.neighbor_list(*b.structure(s), rc, cap, 1e-10, oracle.ALGO_CELLLIST) for s in range(b.n_structs)]


@pytest.mark.gpu
@pytest.mark.parametrize("which", ["direct", "cell"])
@pytest.mark.parametrize("rc,cap", [(8.0, None), (5.0, 12)])
def test_gpu_collate_matches_restatement(gpu_builder, gpu_builder_cell, oracle, syn, which, rc, cap):
    gb = gpu_builder if which == "direct" else gpu_builder_cell
    b = _mixed_batch(syn)
    cap_o = oracle.UNCAPPED if cap is None else cap
    g = gb.build(b.atom_offsets, b.x, b.y, b.z, b.lattices, r_cutoff=rc, max_neighbors=cap_o, rbf_r_cutoff=rc, dr=0.2)
    og = _oracle_graphs(oracle, b, rc, cap_o)
    rng = np.random.default_rng(1)
    keys = rng.integers(0, 5, size=b.n_atoms).astype(np.int32)
    emb = _emb(keys=range(0, 5))
    ek = np.array(sorted(emb), dtype=np.int32)
    et = np.stack([emb[int(k)] for k in ek])
    c = gb.collate(keys, ek, et)
    X, ei, batch = gb.collated_to_host(c)
    per = [keys[b.atom_offsets[s]:b.atom_offsets[s + 1]] for s in range(b.n_structs)]
    Xo, eio, bo = oracle.collate(og, per, emb)
    assert c.n_nodes == b.n_atoms and c.n_edges == eio.shape[1] and c.F == 92 and c.K == g.K
    assert np.array_equal(ei, eio)
    assert np.array_equal(batch, bo)
    assert np.array_equal(X, Xo)                      # same fp64 -> fp32 cast, bit-exact
    assert c.edge_attr == g.rbf                       # edge_attr is the build's feature tensor, not a copy
    info = gb.consumer_info()
    assert info.kernel_launches == 2 and info.bytes_written == 16 * c.n_edges + 8 * c.n_nodes + 4 * 92 * c.n_nodes
    # edge_index / batch only
    c2 = gb.collate()
    _, ei2, b2 = gb.collated_to_host(c2)
    assert c2.F == 0 and not c2.x and np.array_equal(ei2, eio) and np.array_equal(b2, bo)
    # a key the table does not hold: the reference's map::at throws
    bad = keys.copy()
    bad[b.n_atoms // 2] = 9
    with pytest.raises(Exception) as ei_:
        gb.collate(bad, ek, et)
    assert "key not in the embedding table" in str(ei_.value) and f"node {b.n_atoms // 2}" in str(ei_.value)


@pytest.mark.gpu
@pytest.mark.parametrize("F", [1, 7, 8, 92])
def test_gpu_collate_feature_widths(gpu_builder, oracle, syn, F):
    """x gather: widths that are not a multiple of four take the scalar kernel, the others the float4 one; odd edge
    totals shift the 16-byte grid of edge_index row 1 (paired stores peel head and tail per row)."""
    gb = gpu_builder
    b = _mixed_batch(syn)
    rng = np.random.default_rng(F)
    keys = rng.integers(0, 5, size=b.n_atoms).astype(np.int32)
    emb = _emb(F=F, keys=range(0, 5))
    ek = np.array(sorted(emb), dtype=np.int32)
    et = np.stack([emb[int(k)] for k in ek])
    per = [keys[b.atom_offsets[s]:b.atom_offsets[s + 1]] for s in range(b.n_structs)]
    parities = set()
    for rc, cap in ((8.0, oracle.UNCAPPED), (5.0, 11), (6.1, 13), (7.3, oracle.UNCAPPED)):
        gb.build(b.atom_offsets, b.x, b.y, b.z, b.lattices, r_cutoff=rc, max_neighbors=cap, rbf_r_cutoff=rc, dr=0.2)
        og = _oracle_graphs(oracle, b, rc, cap)
        c = gb.collate(keys, ek, et)
        X, ei, batch = gb.collated_to_host(c)
        Xo, eio, bo = oracle.collate(og, per, emb)
        assert c.F == F and np.array_equal(X, Xo) and np.array_equal(ei, eio) and np.array_equal(batch, bo)
        parities.add(c.n_edges & 1)
    if F == 92:
        assert parities == {0, 1}, "the cases are meant to cover even and odd edge totals"


@pytest.mark.gpu
@pytest.mark.parametrize("which", ["direct", "cell"])
def test_gpu_edge_views_and_clouds_match_restatement(gpu_builder, gpu_builder_cell, oracle, syn, which):
    gb = gpu_builder if which == "direct" else gpu_builder_cell
    b = _mixed_batch(syn)
    for rc, cap in ((5.0, 12), (8.0, oracle.UNCAPPED)):      # viewer default; topology caller (uncapped)
        g = gb.build(b.atom_offsets, b.x, b.y, b.z, b.lattices, r_cutoff=rc, max_neighbors=cap, skip_rbf=True)
        h = gb.to_host(g)
        src, tgt, dist, disp = gb.edge_views()
        og = _oracle_graphs(oracle, b, rc, cap)
        e0 = 0
        for s, o in enumerate(og):
            x, y, z, L = b.structure(s)
            d = oracle.displacements(x, y, z, L, o)
            so, to, do, po = oracle.viewer_arrays(o, d)
            e1 = e0 + o.n_edges
            assert np.array_equal(src[e0:e1], so) and np.array_equal(tgt[e0:e1], to)
            assert np.array_equal(dist[e0:e1], do)
            assert np.array_equal(disp[3 * e0:3 * e1], po)
            # neighbourhood clouds of this structure's atoms
            a0, a1 = int(b.atom_offsets[s]), int(b.atom_offsets[s + 1])
            cloud, cptr = gb.neighborhood_clouds(a0, a1, h.row_ptr)
            assert cptr[0] == 0 and cptr[-1] == cloud.shape[0] == o.n_edges + (a1 - a0)
            for i in range(a1 - a0):
                ref = oracle.neighborhood_cloud(x, y, z, o, i, d)
                assert np.array_equal(cloud[cptr[i]:cptr[i + 1]], ref), (s, i)
            e0 = e1
        assert e0 == h.n_edges
        # sub-ranges and the empty range
        cloud, cptr = gb.neighborhood_clouds(3, 3, h.row_ptr)
        assert cloud.shape[0] == 0 and cptr[0] == 0
        only = gb.edge_views(displacements=False)
        assert only[3] is None and np.array_equal(only[2], dist)


@pytest.mark.gpu
def test_gpu_collated_batch_dlpack_zero_copy(pkg, oracle, syn):
    """f1: the collated batch goes to torch through DLPack without a copy (same device pointers), with the dtypes of the
    Python consumer (utilities.py:69-71); live exports pin the arena: build / collate refuse to overwrite it."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    gb = pkg.GraphBuilder(0)
    b = _mixed_batch(syn)
    g = gb.build(b.atom_offsets, b.x, b.y, b.z, b.lattices, r_cutoff=8.0, max_neighbors=oracle.UNCAPPED, rbf_r_cutoff=8.0, dr=0.2)
    h = gb.to_host(g)
    keys = (np.arange(b.n_atoms) % 5).astype(np.int32)
    emb = _emb(keys=range(0, 5))
    ek = np.array(sorted(emb), dtype=np.int32)
    et = np.stack([emb[int(k)] for k in ek])
    c = gb.collate(keys, ek, et)
    X, ei, batch = gb.collated_to_host(c)
    prod = gb.collated_dlpack(c)
    t = {k: torch.from_dlpack(v) for k, v in prod.items()}
    assert gb.live_exports() == 4
    assert t["x"].dtype == torch.float32 and t["edge_attr"].dtype == torch.float32
    assert t["edge_index"].dtype == torch.int64 and t["batch"].dtype == torch.int64
    assert tuple(t["x"].shape) == (b.n_atoms, 92) and tuple(t["edge_index"].shape) == (2, h.n_edges)
    assert tuple(t["edge_attr"].shape) == (h.n_edges, 40) and tuple(t["batch"].shape) == (b.n_atoms,)
    assert t["x"].data_ptr() == c.x and t["edge_index"].data_ptr() == c.edge_index          # zero copy
    assert t["edge_attr"].data_ptr() == g.rbf and t["batch"].data_ptr() == c.batch
    assert all(v.device.type == "cuda" and v.device.index == 0 for v in t.values())
    assert np.array_equal(t["x"].cpu().numpy(), X) and np.array_equal(t["edge_index"].cpu().numpy(), ei)
    assert np.array_equal(t["batch"].cpu().numpy(), batch) and np.array_equal(t["edge_attr"].cpu().numpy(), h.rbf)
    # the arena is pinned while consumers hold the tensors
    with pytest.raises(pkg.binding.DgnnError) as e1:
        gb.build(b.atom_offsets, b.x, b.y, b.z, b.lattices, r_cutoff=5.0, max_neighbors=12, skip_rbf=True)
    assert "still borrow" in str(e1.value)
    with pytest.raises(pkg.binding.DgnnError):
        gb.collate()
    del t
    import gc
    gc.collect()
    assert gb.live_exports() == 0
    # an exported-but-never-consumed capsule releases its pin when it dies
    cap = prod["x"].__dlpack__()
    assert gb.live_exports() == 1
    del cap
    gc.collect()
    assert gb.live_exports() == 0
    gb.build(b.atom_offsets, b.x, b.y, b.z, b.lattices, r_cutoff=5.0, max_neighbors=12, skip_rbf=True)   # free again
    # destroying the builder under a live export defers the release to the tensor's deleter
    c = gb.collate()
    keep = torch.from_dlpack(gb.collated_dlpack(c)["batch"])
    gb.close()
    assert int(keep.sum().item()) == int(batch.sum())
    del keep
    gc.collect()


@pytest.mark.gpu
@pytest.mark.parametrize("which", ["direct", "cell"])
def test_gpu_distance_tables_match_restatement(gpu_builder, gpu_builder_cell, oracle, syn, which):
    """f2: the Betti caller's (deg+1)^2 pairwise-distance tables (ripser_wrapper.cpp:60-66), bit-exact in fp64 against
    the numpy restatement of Eigen's operation order; the compressed fp32 lower triangle is what Ripser is handed."""
    gb = gpu_builder if which == "direct" else gpu_builder_cell
    b = _mixed_batch(syn)
    for rc, cap in ((8.0, oracle.UNCAPPED), (5.0, 12)):
        g = gb.build(b.atom_offsets, b.x, b.y, b.z, b.lattices, r_cutoff=rc, max_neighbors=cap, skip_rbf=True)
        og = _oracle_graphs(oracle, b, rc, cap)
        N = b.n_atoms
        full, fptr = gb.distance_tables(0, N, mode=0)
        low, lptr = gb.distance_tables(0, N, mode=1)
        assert fptr[0] == 0 and lptr[0] == 0 and fptr[-1] == full.shape[0] and lptr[-1] == low.shape[0]
        for s, o in enumerate(og):
            x, y, z, L = b.structure(s)
            d = oracle.displacements(x, y, z, L, o)
            a0 = int(b.atom_offsets[s])
            for i in range(len(x)):
                cloud = oracle.neighborhood_cloud(x, y, z, o, i, d)
                m = cloud.shape[0]
                ref = oracle.distance_table(cloud)
                got = full[fptr[a0 + i]:fptr[a0 + i + 1]]
                assert got.shape[0] == m * m, (s, i)
                assert np.array_equal(got.reshape(m, m).T.view(np.int64), ref.view(np.int64)), (s, i)  # column-major
                assert np.array_equal(low[lptr[a0 + i]:lptr[a0 + i + 1]], oracle.ripser_lower_triangle(ref)), (s, i)
        # a sub-range, the empty range, bad arguments
        sub, sptr = gb.distance_tables(2, 5, mode=1)
        assert np.array_equal(sub, low[lptr[2]:lptr[5]]) and np.array_equal(sptr, lptr[2:6] - lptr[2])
        e, eptr = gb.distance_tables(4, 4, mode=0)
        assert e.shape[0] == 0 and eptr[0] == 0
        gb.distance_tables(0, 3, mode=1)
        info = gb.consumer_info()
        assert info.kernel_launches >= 2 and info.device_ms > 0 and info.bytes_written > 0
        with pytest.raises(Exception):
            gb.distance_tables(0, N + 1)
        with pytest.raises(Exception):
            gb.distance_tables(0, 1, mode=2)


@pytest.mark.gpu
def test_gpu_consumers_reject_stale_graphs(gpu_builder, syn, pkg):
    gb = gpu_builder
    b = syn.c2_batch(2)
    g1 = gb.build(b.atom_offsets, b.x, b.y, b.z, b.lattices, r_cutoff=5.0, max_neighbors=12, skip_rbf=True)
    stale = type(g1).from_buffer_copy(g1)
    b2 = syn.c2_batch(1)
    gb.build(b2.atom_offsets, b2.x, b2.y, b2.z, b2.lattices, r_cutoff=5.0, max_neighbors=12, skip_rbf=True)
    with pytest.raises(pkg.DgnnError):
        gb.collate(g=stale)
