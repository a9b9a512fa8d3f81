"""Cost of exact distance ties: the same 2000 perovskite supercells with jitter (no ties) and without (ideal lattice:
every coordination shell is a run of equal distances the truncated sort key cannot order).  gpurun -- python tools/tie_probe.py"""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry  # noqa: E402

pkg = entry.load_package()
syn = pkg.synthetic
UNCAPPED = 0x7fffffff
gb = pkg.GraphBuilder(0)
only = None  # only=<jitter>,<r_c>,<cap or 0>: one case (for profiler captures)
for kv in sys.argv[1:]:
    k, v = kv.split("=")
    if k == "only":
        only = tuple(float(x) for x in v.split(","))
        continue
    gb.set_option(k, float(v))
for jitter in (0.02, 0.0):
    b = syn.concat([syn.perovskite_supercell(4, 4, 2, seed=s, jitter=jitter)[:4] for s in range(2000)])
    d_in = [torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in (b.atom_offsets, b.x, b.y, b.z, b.lattices)]
    if only and only[0] != jitter:
        continue
    for rc, cap in ((8.0, UNCAPPED), (10.0, 20), (10.0, UNCAPPED)):
        if only and (only[1] != rc or (only[2] == 0) != (cap == UNCAPPED)):
            continue
        ms = []
        for _ in range(4):
            g = gb.build(*d_in, r_cutoff=rc, max_neighbors=cap, rbf_r_cutoff=rc, dr=0.25, device_inputs=True)
            i = gb.info()
            ms.append((i.total_ms, i.count_ms, i.fill_ms))
        t = min(ms)
        print(json.dumps({"jitter": jitter, "r_c": rc, "cap": None if cap == UNCAPPED else cap, "edges": int(g.n_edges),
                          "total_ms": round(t[0], 3), "count_ms": round(t[1], 3), "fill_ms": round(t[2], 3),
                          "launches": int(i.kernel_launches), "tie_run_mean": round(float(i.tie_run_mean), 2),
                          "full_tie_sort": int(i.full_tie_sort)}))
