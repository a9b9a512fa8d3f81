import sys, os, numpy as np
sys.path.insert(0, "/root/repo")
import __graft_entry__ as entry
pkg = entry.load_package(); syn = pkg.synthetic
d = np.load("tests/golden/c1_structures.npz")
parts = []
for nm in d["names"]:
    c = d[f"{nm}__cart"]; parts.append((c[:,0].copy(), c[:,1].copy(), c[:,2].copy(), d[f"{nm}__lattice"].copy()))
b = syn.concat(parts*20)
gb = pkg.GraphBuilder(0)
for cap, dr in ((20, 0.1), (2**64-1, 0.1), (20, None)):
    for i in range(3):
        g = gb.build(b.atom_offsets, b.x, b.y, b.z, b.lattices, r_cutoff=10.0, max_neighbors=cap, rbf_r_cutoff=10.0, dr=dr)
    i = gb.info()
    print(cap, dr, "hit_cap", i.hit_cap, "launches", i.kernel_launches, "count", i.count_ms, "fill", i.fill_ms, "rowcap", i.row_cap, "maxdeg", g.max_degree, "E", g.n_edges)
