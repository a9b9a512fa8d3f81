"""C1 x 200 (1 800 shipped-POSCAR structures, reference defaults) alone: for profiler captures.  gpurun -- python tools/c1x200.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402
import torch  # noqa: E402

pkg = entry.load_package()
syn = pkg.synthetic
d = np.load(os.path.join(ROOT, "tests", "golden", "c1_structures.npz"))
parts = []
for nm in d["names"]:
    c = d[f"{nm}__cart"]
    parts.append((c[:, 0].copy(), c[:, 1].copy(), c[:, 2].copy(), d[f"{nm}__lattice"].copy()))
b = syn.concat(parts * 200)
gb = pkg.GraphBuilder(0)
for kv in sys.argv[1:]:
    k, v = kv.split("=")
    gb.set_option(k, float(v))
t = [torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in (b.atom_offsets, b.x, b.y, b.z, b.lattices)]
for _ in range(4):
    g = gb.build(*t, r_cutoff=10.0, max_neighbors=20, epsilon=1e-10, rbf_r_cutoff=10.0, dr=0.1, device_inputs=True)
    i = gb.info()
print(b.n_structs, b.n_atoms, int(g.n_edges), round(i.total_ms, 3), round(i.count_ms, 3), round(i.fill_ms, 3), i.hit_cap, i.kernel_launches)
