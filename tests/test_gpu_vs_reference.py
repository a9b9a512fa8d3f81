"""GPU parity straight against the REFERENCE'S OWN CODE (oracle/_ref/libref_graph.so = the reference's
src/graph/neighbor_list.cpp, edge_features.cpp, crystal_graph.cpp, structure.cpp compiled verbatim; prebuilt in the
container, travels to the GPU box) -- no restated oracle in between.  Calls go through the C ABI (binding.py -> dgnn_build).

Rows are compared as sets of (idx, distance bits, displacement bits) after a canonical sort (the reference's tie order is
unspecified); rows cut by max_neighbors strictly below their k-th distance.  RBF features: <= 1e-5 relative in fp32
against the reference's fp64 gaussian_rbf (north_star tolerance)."""
import os
from types import SimpleNamespace

import numpy as np
import pytest

from test_ref_pin import R, assert_same_lists, _bits  # noqa: F401  (R is a fixture)

RBF_RTOL = 1e-5
pytestmark = pytest.mark.gpu


def _gpu_rows(oracle, h, b, s, e0):
    """Slice structure s out of a HostGraphs as an OracleGraph-shaped record with re-derived displacements."""
    a0, a1 = int(b.atom_offsets[s]), int(b.atom_offsets[s + 1])
    rp = h.row_ptr[a0:a1 + 1] - h.row_ptr[a0]
    e1 = e0 + int(rp[-1])
    x, y, z, L = b.structure(s)
    g = SimpleNamespace(row_ptr=rp, dst=h.dst[e0:e1], image=h.image[e0:e1], dist=h.dist[e0:e1], n_edges=e1 - e0)
    g.disp = oracle.displacements(x, y, z, L, g)
    return g, e1


def _check_batch(R, oracle, gb, pkg, b, rc, cap, dr, eps=1e-10, what=""):
    capv = pkg.UNCAPPED if cap is None else cap
    g = gb.build(b.atom_offsets, b.x, b.y, b.z, b.lattices, r_cutoff=rc, max_neighbors=capv, epsilon=eps,
                 rbf_r_cutoff=rc, dr=dr)
    h = gb.to_host(g)
    e0 = 0
    for s in range(b.n_structs):
        a0, a1 = int(b.atom_offsets[s]), int(b.atom_offsets[s + 1])
        ref = R.neighbor_list(b.frac[a0:a1], b.lattices[s], rc, R.UNCAPPED if cap is None else cap, eps)
        gs, e1 = _gpu_rows(oracle, h, b, s, e0)
        assert_same_lists(ref, gs, cap, f"{what}[{s}]")
        if ref.n_edges and s % 7 == 0:
            want = np.stack([R.gaussian_rbf(d, rc, dr) for d in ref.dist[:256]])
            got = h.rbf[e0:e0 + want.shape[0]].astype(np.float64)
            assert np.max(np.abs(got - want) / want) <= RBF_RTOL, f"{what}[{s}]: RBF"
        e0 = e1
    assert e0 == h.n_edges


def _with_frac(syn, c1_structures):
    return syn.concat([(x, y, z, L, sp, fr) for (x, y, z, L, sp, fr) in c1_structures.values()])


@pytest.mark.parametrize("rc,cap,dr", [(10.0, 20, 0.1), (5.0, 12, 0.2), (8.0, None, 0.2), (10.0, None, 0.1)])
def test_gpu_matches_compiled_reference_on_shipped_poscars(R, oracle, pkg, syn, gpu_builder, c1_structures, rc, cap, dr):
    _check_batch(R, oracle, gpu_builder, pkg, _with_frac(syn, c1_structures), rc, cap, dr, what=f"C1 rc={rc} cap={cap}")


def test_gpu_cell_list_path_matches_compiled_reference(R, oracle, pkg, syn, gpu_builder_cell, c1_structures):
    _check_batch(R, oracle, gpu_builder_cell, pkg, _with_frac(syn, c1_structures), 8.0, None, 0.2, what="C1 cell-list")
    _check_batch(R, oracle, gpu_builder_cell, pkg, _with_frac(syn, c1_structures), 10.0, 20, 0.1, what="C1 cell-list cap")


def test_gpu_matches_compiled_reference_on_synthetic_configs(R, oracle, pkg, syn, gpu_builder):
    _check_batch(R, oracle, gpu_builder, pkg, syn.c2_batch(48, seed0=1234), 8.0, None, 0.2, what="C2")
    _check_batch(R, oracle, gpu_builder, pkg, syn.c2_batch(24, seed0=77), 8.0, 20, 0.2, what="C2 cap 20")
    _check_batch(R, oracle, gpu_builder, pkg, syn.c3_batch(150, seed0=777), 8.0, None, 0.2, what="C3")
    _check_batch(R, oracle, gpu_builder, pkg, syn.c3_batch(60, seed0=31), 8.0, 20, 0.2, what="C3 cap 20")
    adv = syn.adversarial_cells()
    _check_batch(R, oracle, gpu_builder, pkg, adv, 8.0, None, 0.2, what="adversarial")
    _check_batch(R, oracle, gpu_builder, pkg, adv, 6.0, None, 0.2, eps=0.0, what="adversarial eps=0")
    _check_batch(R, oracle, gpu_builder, pkg, adv, 6.0, None, 0.2, eps=6.1, what="adversarial eps=6.1")
