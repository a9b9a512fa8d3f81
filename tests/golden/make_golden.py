#!/usr/bin/env python
"""Generate the committed golden fixtures under tests/golden/.

Run in the build container only (it reads the reference's shipped POSCAR *data* files, which do not exist on the
GPU box):   python tests/golden/make_golden.py

Outputs
  c1_structures.npz   the 9 shipped POSCARs (reference web/public/data/structures/*.vasp) parsed to
                      lattice / fractional / Cartesian arrays (config C1 inputs)
  c1_expected.json    per file and per parameter set: edge count, degree histogram digest, SHA-256 of the canonical
                      edge list (dst, image, distance bits) from the exhaustive-scan oracle, n_near_cutoff, tie rows;
                      plus the independent expected-E table of SURVEY.md Appendix B (numpy/scipy probe) they must equal
  c1_capped_edges.npz full oracle edge lists for the reference-default parameters (r_c=10, cap 20) -- small
  kat.json            analytic known answers (simple-cubic shells, closed-form Gaussians)

The reference has no tests / golden vectors for this path.  The fixtures are therefore generated
from the REFERENCE'S OWN CODE run here: oracle/_ref/libref_graph.so = the reference's src/graph/*.cpp, structure.cpp and
vasp_parser.cpp compiled verbatim (oracle/Makefile `ref`).  Inputs come from the reference's io::parse_vasp, positions
from crystal::Structure, edge counts / distance sums from graph::NeighborList; the canonical edge-list digest (which
needs the integer image offsets the reference does not store) comes from the oracle AFTER the oracle's rows have been
checked to be the reference's rows as sets of (idx, distance bits, displacement bits) -- tests/test_ref_pin.py keeps
checking exactly that.  Any later change to oracle/ that alters a digest fails tests/test_oracle.py.
"""
import glob
import hashlib
import importlib.util
import json
import math
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import oracle_py as O  # noqa: E402
import ref_py as R  # noqa: E402
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)
from test_ref_pin import assert_same_lists  # noqa: E402

spec = importlib.util.spec_from_file_location("dgnn_synthetic", os.path.join(ROOT, "defect-gnn-cpp_b200", "synthetic.py"))
syn = importlib.util.module_from_spec(spec)
sys.modules["dgnn_synthetic"] = syn
spec.loader.exec_module(syn)

REF_DIR = "/root/reference/web/public/data/structures"

# (name, r_cutoff, max_neighbors, dr)  -- SURVEY.md section 8c(i)
PARAM_SETS = [
    ("reference_default", 10.0, 20, 0.1),     # neighbor_list.hpp:19-22, crystal_graph.hpp:14-19
    ("viewer_default", 5.0, 12, None),        # web/src/app/visualizer/page.tsx:26-27
    ("benchmark_uncapped", 8.0, O.UNCAPPED, 0.2),
    ("topology_uncapped", 10.0, O.UNCAPPED, 0.1),  # betti_features.cpp:107
]

# E at r_c = 10 / 8 / 5, uncapped (independent numpy/scipy probe)
SURVEY_E = {
    "1": (64872, 32868, 8748), "1_1": (64156, 32510, 8648), "1_2": (64150, 32510, 8656),
    "741": (28992, 14184, 3504), "741_1": (28502, 13942, 3444), "741_2": (28502, 13942, 3440),
    "1046": (52812, 28728, 6750), "1046_1": (51996, 28312, 6650), "1046_2": (52056, 28336, 6662),
}


def edge_digest(g) -> str:
    h = hashlib.sha256()
    h.update(np.ascontiguousarray(g.row_ptr, dtype=np.int64).tobytes())
    h.update(np.ascontiguousarray(g.dst, dtype=np.int32).tobytes())
    h.update(np.ascontiguousarray(g.image, dtype=np.int8).tobytes())
    h.update(np.ascontiguousarray(g.dist, dtype=np.float64).tobytes())
    return h.hexdigest()


def main():
    files = sorted(glob.glob(os.path.join(REF_DIR, "*.vasp")))
    assert len(files) == 9, files
    names, store = [], {}
    expected = {}
    capped = {}
    for f in files:
        name = os.path.splitext(os.path.basename(f))[0]
        L, frac, species = R.parse_vasp(f)                       # the reference's parser
        Lp, fp, sp_ = syn.parse_poscar(open(f).read())           # the fixture loader must agree with it bit for bit
        assert np.array_equal(L, Lp) and np.array_equal(frac, fp) and np.array_equal(species, sp_), name
        st = R.Structure(frac, L, species, 3.0, 1)
        pos = st.neighbor_list().pos                             # crystal::Structure positions (structure.cpp:17)
        st.close()
        x, y, z = (np.ascontiguousarray(pos[:, k]) for k in range(3))
        assert all(np.array_equal(a, b) for a, b in zip((x, y, z), syn.frac_to_cart(frac, L))), name
        names.append(name)
        store[f"{name}__lattice"] = L
        store[f"{name}__frac"] = frac
        store[f"{name}__cart"] = np.stack([x, y, z], axis=1)
        store[f"{name}__species"] = species
        expected[name] = {"n_atoms": int(len(x))}
        for pname, rc, cap, dr in PARAM_SETS:
            g = O.neighbor_list(x, y, z, L, rc, cap, 1e-10, O.ALGO_BRUTE)
            ref = R.neighbor_list(frac, L, rc, cap, 1e-10)
            assert_same_lists(ref, g, None if cap == O.UNCAPPED else cap, f"{name}/{pname}")
            assert ref.n_edges == g.n_edges and float(np.sum(ref.dist)) == float(np.sum(g.dist))
            deg = np.diff(g.row_ptr)
            rec = {"source": "oracle/_ref",
                "r_cutoff": rc, "max_neighbors": (None if cap == O.UNCAPPED else cap), "dr": dr,
                "num_images": g.num_images, "n_edges": g.n_edges, "deg_min": int(deg.min()), "deg_max": int(deg.max()),
                "n_near_cutoff": g.n_near_cutoff, "n_tie_rows": g.n_tie_rows, "edge_sha256": edge_digest(g),
                "dist_sum": float(np.sum(g.dist)),
            }
            if dr is not None:
                feats = O.edge_attr(g.dist[: min(g.n_edges, 4096)], rc, dr)
                rec["K"] = int(feats.shape[1])
                rec["rbf_sum_first4096"] = float(feats.sum())
            expected[name][pname] = rec
            if pname == "reference_default":
                capped[f"{name}__row_ptr"] = g.row_ptr
                capped[f"{name}__dst"] = g.dst
                capped[f"{name}__image"] = g.image
                capped[f"{name}__dist"] = g.dist
        # cross-check against the independent SURVEY Appendix-B probe
        for rc, e_survey in zip((10.0, 8.0, 5.0), SURVEY_E[name]):
            e = O.neighbor_list(x, y, z, L, rc, O.UNCAPPED, 1e-10, O.ALGO_CELLLIST).n_edges
            assert e == e_survey, (name, rc, e, e_survey)
        expected[name]["survey_appendix_b_E_rc10_8_5"] = list(SURVEY_E[name])
        print(name, expected[name]["n_atoms"], {p[0]: expected[name][p[0]]["n_edges"] for p in PARAM_SETS})
    store["names"] = np.array(names)
    np.savez_compressed(os.path.join(HERE, "c1_structures.npz"), **store)
    np.savez_compressed(os.path.join(HERE, "c1_capped_edges.npz"), **capped)
    with open(os.path.join(HERE, "c1_expected.json"), "w") as fh:
        json.dump(expected, fh, indent=1, sort_keys=True)

    # ---- analytic known answers -------------------------------------------------------------------------------
    kat = {"simple_cubic_shells": {"a": 1.0, "r_cutoff": 3.05,
                                   "d2_to_multiplicity": {"1": 6, "2": 12, "3": 8, "4": 6, "5": 24, "6": 24, "8": 12, "9": 30}}}
    rbf = []
    for rc, dr in ((8.0, 0.2), (10.0, 0.1)):
        sigma = rc / 3.0
        norm = 1.0 / (sigma * math.sqrt(2.0 * math.pi))
        for d in (0.0, rc / 2.0, rc - 1e-9):
            K = int(math.floor(rc / dr))
            vals = [norm * math.exp(-0.5 * ((k * dr - d) ** 2) / (sigma * sigma)) for k in range(K)]
            rbf.append({"r_cutoff": rc, "dr": dr, "distance": d, "K": K, "values": vals})
    kat["gaussian_rbf_closed_form"] = rbf
    # C2 sample structure summary (seed 1234): pins the generator + oracle together
    x, y, z, L, sp = syn.perovskite_supercell(4, 4, 2, seed=1234)
    g = O.neighbor_list(x, y, z, L, 8.0, O.UNCAPPED, 1e-10, O.ALGO_BRUTE)
    kat["c2_seed1234"] = {"n_atoms": int(len(x)), "n_edges": g.n_edges, "n_near_cutoff": g.n_near_cutoff,
                          "edge_sha256": edge_digest(g), "pos_sha256": hashlib.sha256(
                              np.stack([x, y, z]).tobytes()).hexdigest()}
    with open(os.path.join(HERE, "kat.json"), "w") as fh:
        json.dump(kat, fh, indent=1, sort_keys=True)
    print("wrote golden fixtures to", HERE)


if __name__ == "__main__":
    main()
