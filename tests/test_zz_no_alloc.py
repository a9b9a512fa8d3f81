"""dgnn_row_offsets_noalloc (include/dgnn_cabi.h): the allocation-free variant of the degree scan, checked through the
C ABI against the existing entry point's output -- dgnn_build's row_ptr.  No counterpart in the reference;
collected last on purpose (tests/test_zz_*), so that a failure here cannot hide another
test under `pytest -x`.  Integer work: the comparison is bit-exact."""
import numpy as np
import pytest


def _device_row_ptr_and_degrees(torch, gb, b, rc, cap):
    """dgnn_build (the existing entry point) on `b`: its row_ptr and the per-row degrees, both on the device."""
    g = gb.build(b.atom_offsets, b.x, b.y, b.z, b.lattices, r_cutoff=rc, max_neighbors=cap, epsilon=1e-10,
                 rbf_r_cutoff=rc, dr=0.2)
    # row_ptr of the build, copied out of the ctx arena with the library's own copy entry point
    h_rp = np.ascontiguousarray(gb.to_host(g, with_rbf=False).row_ptr, dtype=np.int64)
    assert h_rp.shape[0] == int(b.n_atoms) + 1
    assert int(h_rp[-1]) == int(g.n_edges)
    deg = torch.from_numpy(np.diff(h_rp).astype(np.int32)).cuda()
    return h_rp, deg


@pytest.mark.gpu
def test_row_offsets_noalloc_matches_build_twice_in_a_row(pkg, syn):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    binding = pkg.binding
    gb = pkg.GraphBuilder(0)
    # two different inputs back to back: the second call finds the first one's tile sums in the shared scratch
    cases = [(syn.c2_batch(40, seed0=5), 8.0, pkg.UNCAPPED),           # 6360 rows: 4 tiles, ragged tail
             (syn.c3_batch(60, seed0=3), 8.0, 20)]                     # a few hundred rows: 1 tile
    outs = []
    for b, rc, cap in cases:
        ref_rp, deg = _device_row_ptr_and_degrees(torch, gb, b, rc, cap)
        out = torch.full((deg.numel() + 1,), -1, dtype=torch.int64, device="cuda")
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        binding.row_offsets_noalloc(0, deg.data_ptr(), deg.numel(), out.data_ptr(),
                                    torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before
        got = out.cpu().numpy()
        assert np.array_equal(got, ref_rp), f"row offsets differ from dgnn_build's ({deg.numel()} rows)"
        outs.append((deg, ref_rp))
    # ... and the first input again (scratch now holds the smaller second call's sums), default stream
    deg, ref_rp = outs[0]
    out = torch.full((deg.numel() + 1,), -1, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    binding.row_offsets_noalloc(0, deg.data_ptr(), deg.numel(), out.data_ptr(), None)
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy(), ref_rp)
    gb.close()


@pytest.mark.gpu
def test_row_offsets_noalloc_sizes_and_limit(pkg):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    binding = pkg.binding
    rng = np.random.default_rng(0)
    # empty input, one row, tile boundaries, many tiles (more than one round of the 1024-wide sum scan), int64 totals
    for n in (0, 1, 2047, 2048, 2049, 3_000_000):
        d = rng.integers(0, 400, size=n, dtype=np.int32)
        if n == 3_000_000:
            d[:] = 2_000                                                # total 6e9 > 2^32
        deg = torch.from_numpy(d).cuda()
        out = torch.full((n + 1,), -1, dtype=torch.int64, device="cuda")
        for _ in range(2):
            binding.row_offsets_noalloc(0, deg.data_ptr() if n else 0, n, out.data_ptr(), None)
            torch.cuda.synchronize()
            want = np.concatenate([[0], np.cumsum(d.astype(np.int64))])
            assert np.array_equal(out.cpu().numpy(), want), n
    # beyond the fixed scratch: rejected before anything is launched, with a message that names the limit
    out = torch.zeros(2, dtype=torch.int64, device="cuda")
    with pytest.raises(pkg.DgnnError) as ei:
        binding.row_offsets_noalloc(0, out.data_ptr(), binding.NOALLOC_MAX_ROWS + 1, out.data_ptr(), None)
    assert ei.value.code == -9 and "DGNN_NOALLOC_MAX_ROWS" in str(ei.value)
    with pytest.raises(pkg.DgnnError) as ei:
        binding.row_offsets_noalloc(0, 0, 5, out.data_ptr(), None)
    assert ei.value.code == -1
    assert int(out.sum()) == 0
