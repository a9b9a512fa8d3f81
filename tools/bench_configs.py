#!/usr/bin/env python
"""Throughput of the non-headline BASELINE.json configs on one B200 (C1 shipped POSCARs, C3 triclinic short-vector
cells, C4 one 1M-atom cell, C2 capped) -- one JSON line per config, device-resident inputs, CUDA-event phase times from
dgnn_build_info.  The headline config (C2 uncapped) is bench.py's job.  Usage: python tools/bench_configs.py [--reps 5]
"""
import argparse
import json
import os
import statistics
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402

UNCAPPED = 2**64 - 1


def run(gb, torch, name, b, rc, cap, dr, reps, peak):
    dev = torch.device("cuda:0")
    t = [torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in (b.atom_offsets, b.x, b.y, b.z, b.lattices)]
    torch.cuda.synchronize()
    infos = []
    for i in range(reps + 2):
        g = gb.build(*t, r_cutoff=rc, max_neighbors=cap, epsilon=1e-10, rbf_r_cutoff=rc, dr=dr, device_inputs=True)
        if i >= 2:
            infos.append(gb.info())
    E, K = int(g.n_edges), int(g.K)
    ms = statistics.median(i.total_ms for i in infos)
    bytes_all = sum(int(x.numel() * x.element_size()) for x in t) + 8 * (b.n_atoms + 1) + E * (16 + 4 * K)
    line = {"config": name, "n_structs": b.n_structs, "n_atoms": b.n_atoms, "n_edges": E, "K": K,
            "r_cutoff": rc, "max_neighbors": "uncapped" if cap == UNCAPPED else cap,
            "ms_per_build": ms, "edges_per_sec": E / (ms * 1e-3), "structures_per_sec": b.n_structs / (ms * 1e-3),
            "whole_path_gbs": bytes_all / (ms * 1e-3) / 1e9, "frac_of_measured_peak": bytes_all / (ms * 1e-3) / 1e9 / peak,
            "phases_ms": {k: statistics.median(getattr(i, k) for i in infos)
                          for k in ("bin_ms", "count_ms", "fill_ms", "rbf_ms")},
            "launches": infos[-1].kernel_launches, "onepass": infos[-1].onepass, "onepass_fallback_rows": infos[-1].onepass_fallback_rows, "hit_cap": infos[-1].hit_cap, "cell_list_structs": infos[-1].n_cell_structs,
            "max_degree": int(g.max_degree), "near_cutoff_pairs": int(g.n_near_cutoff), "tie_rows": int(g.n_tie_rows)}
    print(json.dumps(line), flush=True)
    del t
    return line


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--only", default="")
    ap.add_argument("--opt", action="append", default=[], help="name=value for dgnn_set_option")
    args = ap.parse_args()
    import torch
    pkg = entry.load_package()
    syn = pkg.synthetic
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        peak = 6650.0
    gb = pkg.GraphBuilder(0)
    for kv in args.opt:
        k, v = kv.split("=")
        gb.set_option(k, float(v))
    want = set(args.only.split(",")) if args.only else None
    def on(n):
        return want is None or n in want
    if on("C1"):
        d = np.load(os.path.join(ROOT, "tests", "golden", "c1_structures.npz"))
        parts = []
        for nm in d["names"]:
            c = d[f"{nm}__cart"]
            parts.append((c[:, 0].copy(), c[:, 1].copy(), c[:, 2].copy(), d[f"{nm}__lattice"].copy()))
        b1 = syn.concat(parts)
        run(gb, torch, "C1 9 shipped POSCARs, reference defaults r_c=10 cap=20 dr=0.1", b1, 10.0, 20, 0.1, args.reps, peak)
        run(gb, torch, "C1 9 shipped POSCARs, topology caller r_c=10 uncapped dr=0.1", b1, 10.0, UNCAPPED, 0.1, args.reps, peak)
        # the reference API's own shape: one structure per call (NeighborList + CrystalGraph constructors); wall clock
        # per call through the C ABI with host inputs, H2D and all syncs included
        import time
        singles = [syn.concat([p]) for p in parts]
        for onepass in (-1, 0):
            gb.set_option("onepass", onepass)
            for _ in range(3):
                for b in singles:
                    gb.build(b.atom_offsets, b.x, b.y, b.z, b.lattices, r_cutoff=10.0, max_neighbors=20, epsilon=1e-10,
                             rbf_r_cutoff=10.0, dr=0.1)
            t0 = time.perf_counter()
            reps = 20
            for _ in range(reps):
                for b in singles:
                    gb.build(b.atom_offsets, b.x, b.y, b.z, b.lattices, r_cutoff=10.0, max_neighbors=20, epsilon=1e-10,
                             rbf_r_cutoff=10.0, dr=0.1)
            dt = (time.perf_counter() - t0) / (reps * len(singles))
            print(json.dumps({"config": "C1 one POSCAR per dgnn_build call (reference API shape), host inputs, wall clock",
                              "onepass_option": onepass, "onepass": gb.info().onepass, "ms_per_call": dt * 1e3,
                              "structures_per_sec": 1.0 / dt, "launches": gb.info().kernel_launches}), flush=True)
        gb.set_option("onepass", -1)
        for kv in args.opt:
            k, v = kv.split("=")
            gb.set_option(k, float(v))
        big = syn.concat(parts * 200)
        run(gb, torch, "C1 x200 (1800 POSCAR structures), reference defaults", big, 10.0, 20, 0.1, args.reps, peak)
    if on("C2cap"):
        b2 = syn.c2_batch(10_000)
        run(gb, torch, "C2 10k perovskite supercells, cap=20 (secondary)", b2, 8.0, 20, 0.2, args.reps, peak)
    if on("F1"):
        # consumers of the CSR: device time of the kernels, algorithmic bytes / time against the copy peak
        b2 = syn.c2_batch(10_000)
        dev = torch.device("cuda:0")
        t = [torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in (b2.atom_offsets, b2.x, b2.y, b2.z, b2.lattices)]
        g = gb.build(*t, r_cutoff=8.0, max_neighbors=UNCAPPED, epsilon=1e-10, rbf_r_cutoff=8.0, dr=0.2, device_inputs=True)
        rng = np.random.default_rng(0)
        keys = torch.from_numpy(rng.integers(1, 101, size=b2.n_atoms).astype(np.int32)).to(dev)
        ek = np.arange(1, 101, dtype=np.int32)
        et = rng.standard_normal((100, 92))
        ms, info = [], None
        for i in range(args.reps + 2):
            c = gb.collate(keys, ek, et, device_keys=True)
            info = gb.consumer_info()
            if i >= 2:
                ms.append(info.device_ms)
        m = statistics.median(ms)
        tot = info.bytes_read + info.bytes_written
        print(json.dumps({"config": "F1 dgnn_collate on C2 10k uncapped (x [N,92] f32 gather, edge_index [2,E] i64, batch [N] i64)",
                          "n_nodes": int(c.n_nodes), "n_edges": int(c.n_edges), "device_ms": m,
                          "algorithmic_bytes": int(tot), "achieved_gbs": tot / (m * 1e-3) / 1e9,
                          "frac_of_measured_peak": tot / (m * 1e-3) / 1e9 / peak, "launches": info.kernel_launches}), flush=True)
        del t
        b2c = syn.c2_batch(2_000)
        g = gb.build(b2c.atom_offsets, b2c.x, b2c.y, b2c.z, b2c.lattices, r_cutoff=5.0, max_neighbors=12, skip_rbf=True)
        ms = []
        for i in range(args.reps + 2):
            gb.edge_views()
            info = gb.consumer_info()
            if i >= 2:
                ms.append(info.device_ms)
        m = statistics.median(ms)
        tot = info.bytes_read + info.bytes_written
        print(json.dumps({"config": "F4 dgnn_edge_views on 2k C2 structures, viewer params r_c=5 cap=12 (kernel only; D2H outside)",
                          "n_edges": int(g.n_edges), "device_ms": m, "algorithmic_bytes": int(tot),
                          "achieved_gbs": tot / (m * 1e-3) / 1e9, "launches": info.kernel_launches}), flush=True)
    if on("C3"):
        b3 = syn.c3_batch(10_000)
        run(gb, torch, "C3 10k triclinic short-vector cells r_c=8 K=40 uncapped", b3, 8.0, UNCAPPED, 0.2, args.reps, peak)
    if on("C4"):
        b4 = syn.c4_structure()
        run(gb, torch, "C4 one 1,009,489-atom supercell r_c=8 K=40 uncapped", b4, 8.0, UNCAPPED, 0.2, args.reps, peak)
    gb.close()


if __name__ == "__main__":
    main()
