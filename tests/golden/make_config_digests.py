#!/usr/bin/env python
"""Edge totals of the BASELINE configs' synthetic workloads from the CPU oracle (pinned on oracle/_ref by
tests/test_ref_pin.py) -> tests/golden/config_digests.json.  bench.py compares the GPU builds of its `configs` record
with these committed numbers (it may not execute oracle/ outside the cpu_baseline leg).  Run in the build container:
    python tests/golden/make_config_digests.py        (C4: one 1M-atom cell, ~8 GB of host memory, a few minutes)"""
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402

O = entry.load_oracle()
syn = entry.load_package().synthetic
out = {}


def rec(name, b, rc, cap, algo=None):
    t0 = time.time()
    counts = O.batch_edge_counts(b.atom_offsets, b.x, b.y, b.z, b.lattices, rc, O.UNCAPPED if cap is None else cap,
                                 1e-10, O.ALGO_CELLLIST if algo is None else algo, 0)
    out[name] = {"n_structs": b.n_structs, "n_atoms": b.n_atoms, "r_cutoff": rc, "max_neighbors": cap,
                 "n_edges": int(counts.sum()), "edges_first_structs": [int(c) for c in counts[:8]],
                 "source": "oracle/dgnn_oracle.cpp (cell-list algorithm), pinned on oracle/_ref"}
    print(name, out[name]["n_edges"], f"{time.time() - t0:.1f} s", flush=True)


d = np.load(os.path.join(HERE, "c1_structures.npz"))
parts = []
for nm in d["names"]:
    c = d[f"{nm}__cart"]
    parts.append((c[:, 0].copy(), c[:, 1].copy(), c[:, 2].copy(), d[f"{nm}__lattice"].copy()))
rec("C1_reference_defaults", syn.concat(parts), 10.0, 20)
rec("C1_uncapped_rc10", syn.concat(parts), 10.0, None)
rec("C2_10k_uncapped", syn.c2_batch(10_000), 8.0, None)
rec("C2_10k_cap20", syn.c2_batch(10_000), 8.0, 20)
rec("C3_10k_uncapped", syn.c3_batch(10_000), 8.0, None)
if "--no-c4" not in sys.argv:
    rec("C4_1M_atom_cell", syn.c4_structure(), 8.0, None)
json.dump(out, open(os.path.join(HERE, "config_digests.json"), "w"), indent=1, sort_keys=True)
print("wrote", os.path.join(HERE, "config_digests.json"))
