"""dgnn_build_stream (BASELINE config 5: jobs larger than one edge arena, built chunk by chunk through two arenas while
the consumer works on the previous chunk): every chunk must equal the one-shot dgnn_build of the same structures bit for
bit, chunks arrive in order, totals add up, callback failures surface as errors."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_streamed_build_equals_chunkwise_builds(pkg, oracle, syn, gpu_builder):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    job = syn.concat([p for s in range(11) for p in (syn.perovskite_supercell(4, 4, 2, seed=900 + s)[:4],
                                                      syn.triclinic_cell(40 + s)[:4])])
    bounds = [(0, 5), (5, 6), (6, 6), (6, 14), (14, 22)]     # ragged chunks, one of them empty
    gs = pkg.GraphBuilder(0)
    gs.set_option("onepass", 0)                              # the option log is replayed on the twin arena
    got, order = {}, []

    def source(c):
        if c >= len(bounds):
            return None
        b = job.slice(*bounds[c])
        return b.atom_offsets, b.x, b.y, b.z, b.lattices

    def sink(c, g, arena):
        order.append(c)
        got[c] = arena.to_host(g)       # the arena that owns this chunk (the other one is building chunk c + 1)

    info = gs.build_stream(source, sink, r_cutoff=8.0, max_neighbors=pkg.UNCAPPED, rbf_r_cutoff=8.0, dr=0.2)
    assert order == list(range(len(bounds))) and info.n_chunks == len(bounds)
    assert info.n_structs == job.n_structs and info.n_atoms == job.n_atoms
    tot = 0
    for c, (s0, s1) in enumerate(bounds):
        b = job.slice(s0, s1)
        g = gpu_builder.build(b.atom_offsets, b.x, b.y, b.z, b.lattices, r_cutoff=8.0, max_neighbors=pkg.UNCAPPED,
                              rbf_r_cutoff=8.0, dr=0.2)
        ref = gpu_builder.to_host(g)
        for f in ("row_ptr", "dst", "image", "dist", "rbf", "struct_edge_offsets"):
            assert np.array_equal(getattr(got[c], f), getattr(ref, f)), (c, f)
        tot += ref.n_edges
    assert info.n_edges == tot and info.wall_ms > 0 and info.build_device_ms > 0
    # ... and against the CPU oracle for one chunk
    b = job.slice(*bounds[3])
    e0 = 0
    for s in range(b.n_structs):
        x, y, z, L = b.structure(s)
        o = oracle.neighbor_list(x, y, z, L, 8.0, oracle.UNCAPPED, 1e-10, oracle.ALGO_CELLLIST)
        assert np.array_equal(got[3].dst[e0:e0 + o.n_edges], o.dst) and np.array_equal(
            got[3].dist[e0:e0 + o.n_edges].view(np.int64), o.dist.view(np.int64))
        e0 += o.n_edges
    # callback failures: no hang, the error reaches the caller
    with pytest.raises(ValueError):
        gs.build_stream(lambda c: (_ for _ in ()).throw(ValueError("source broke")), None, r_cutoff=8.0)

    def bad_sink(c, g, arena):
        raise KeyError("sink broke")
    with pytest.raises(KeyError):
        gs.build_stream(source, bad_sink, r_cutoff=8.0, max_neighbors=pkg.UNCAPPED, rbf_r_cutoff=8.0, dr=0.2)
    # the context is still usable afterwards
    info2 = gs.build_stream(source, None, r_cutoff=8.0, max_neighbors=20, rbf_r_cutoff=8.0, dr=0.2)
    assert info2.n_chunks == len(bounds)
    gs.close()
