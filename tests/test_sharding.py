"""CPU tests of the structure-sharded multi-GPU host logic (no collective on the data path): partitioning,
host-side gather of per-shard edge counts over torch.distributed (gloo, world_size 2), global offset scan."""
import os
import socket

import numpy as np
import pytest


def test_partition_covers_everything(pkg):
    sh = pkg.sharding
    rng = np.random.default_rng(0)
    for s in (0, 1, 5, 1000):
        for w in (1, 2, 4, 8):
            weights = rng.uniform(0.5, 2.0, size=s)
            parts = sh.partition_structures(weights, w)
            assert len(parts) == w and parts[0][0] == 0 and parts[-1][1] == s
            for (a, b), (c, d) in zip(parts[:-1], parts[1:]):
                assert b == c and a <= b
            if s >= 100:
                loads = np.array([weights[a:b].sum() for a, b in parts])
                assert loads.max() / loads.mean() < 1.05


def test_partition_balances_by_predicted_edges(pkg, syn):
    sh = pkg.sharding
    small = syn.c2_batch(6, dims=(2, 2, 2))
    big = syn.c2_batch(2, dims=(4, 4, 4))
    b = syn.concat([small.structure(s) for s in range(6)] + [big.structure(s) for s in range(2)])
    w = sh.predicted_edges(b.atom_offsets, b.lattices, 8.0)
    assert w[-1] > 6 * w[0]
    parts = sh.partition_structures(w, 2)
    assert parts[0][1] >= 6  # the six small cells weigh less than one big one


def test_global_offsets_scan(pkg):
    sh = pkg.sharding
    offs = sh.global_edge_offsets([np.array([3, 0, 2]), np.array([], dtype=np.int64), np.array([5])])
    assert offs.tolist() == [0, 3, 3, 5, 10]
    assert sh.shard_edge_base([np.array([3, 0, 2]), np.array([5])]).tolist() == [0, 5, 10]


def _worker(rank, world, port, q):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    import torch.distributed as dist
    import __graft_entry__ as entry
    pkg = entry.load_package()
    O = entry.load_oracle()
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        syn, sh = pkg.synthetic, pkg.sharding
        b = syn.c3_batch(9, seed0=4242)  # every rank builds the same global batch description
        parts = sh.partition_structures(sh.predicted_edges(b.atom_offsets, b.lattices, 8.0), world)
        s0, s1 = parts[rank]
        mine = b.slice(s0, s1)
        # the per-shard build is the GPU's job; here the oracle stands in for it (CPU test of the host logic only)
        counts = O.batch_edge_counts(mine.atom_offsets, mine.x, mine.y, mine.z, mine.lattices, 8.0, threads=1)
        gathered = sh.gather_edge_counts(counts)
        offs = sh.global_edge_offsets(gathered)
        q.put((rank, parts, offs.tolist()))
    finally:
        dist.destroy_process_group()


def test_gloo_world2_gather_and_offsets(pkg, oracle, syn):
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    b = syn.c3_batch(9, seed0=4242)
    ref = oracle.batch_edge_counts(b.atom_offsets, b.x, b.y, b.z, b.lattices, 8.0, threads=1)
    expect = np.concatenate([[0], np.cumsum(ref)]).tolist()
    for rank, parts, offs in res:
        assert offs == expect            # both ranks derive the same global offsets
        assert parts[0][1] == parts[1][0]


def test_stream_plan_covers_the_job_once(pkg):
    """Config 5 sharding: every structure of the job is in exactly one chunk of exactly one rank, in order."""
    sh = pkg.sharding
    for total, world, chunk in ((1_000_000, 8, 10_000), (1_000_003, 4, 12_345), (17, 8, 5), (0, 2, 10), (5, 8, 3)):
        seen = []
        for r in range(world):
            plan = sh.stream_plan(total, world, r, chunk)
            assert all(0 < b - a <= chunk for a, b in plan)
            assert all(plan[i][1] == plan[i + 1][0] for i in range(len(plan) - 1))
            seen += plan
        flat = [s for a, b in seen for s in (a, b)]
        assert flat == sorted(flat)
        assert sum(b - a for a, b in seen) == total and (not seen or (seen[0][0] == 0 and seen[-1][1] == total))
    with pytest.raises(ValueError):
        sh.stream_plan(10, 0, 0, 5)
