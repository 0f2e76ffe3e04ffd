"""world_size-2 gloo test of the N>1 path's host logic: contiguous shards of the configuration
batch, no data-path collective, results concatenated (all_gather of the fixed-width interval table
is the optional exchange step)."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT


def _worker(rank, world, port, pack, q, out_path):
    sys.path.insert(0, ROOT)
    import __graft_entry__ as g

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    bench = g._load("sirrt_bench", os.path.join(ROOT, "bench.py"))
    N = q.shape[1]
    n0, n1 = bench.shard_range(N, rank, world)          # bench.py's own partition of ONE batch ...
    o = g.load_oracle().Oracle(pack)  # stands in for the per-rank GPU scene on this CPU-only box
    table, _ = o.safe_intervals_batch(np.ascontiguousarray(q[:, n0:n1]), 8)
    full = bench.gather_table_shards(torch.from_numpy(table), N, rank, world, dist)   # ... and its all-gather of the table shards
    assert full.shape == (N, 17)
    total = torch.tensor([n1 - n0], dtype=torch.int64)
    dist.all_reduce(total)
    assert int(total) == N
    np.save(out_path + f".{rank}.npy", full.numpy())   # EVERY rank must hold all rows
    dist.destroy_process_group()


@pytest.mark.parametrize("n,world,port", [(64, 2, 29531), (65, 2, 29535), (67, 3, 29537)])
def test_bench_partition_reassembles_on_every_rank(tmp_path, sp, orc, packs, n, world, port):
    """bench.py's N>1 data path on CPU (gloo): contiguous blocks of ONE batch, interval-table shards all-gathered; equal
    blocks, blocks that differ by one row (padding path), three ranks"""
    pk = packs("synA_8_50")
    q = sp.random_configs(sp.read_pack(pk).robot, n, 3)
    out = str(tmp_path / "gathered")
    mp.spawn(_worker, args=(world, port, pk, q, out), nprocs=world, join=True)
    full, _ = orc.Oracle(pk).safe_intervals_batch(q, 8)
    for r in range(world):
        assert np.array_equal(np.load(out + f".{r}.npy"), full), r


def test_shard_range_covers_everything():
    sys.path.insert(0, ROOT)
    import __graft_entry__ as g

    bench = g._load("sirrt_bench", os.path.join(ROOT, "bench.py"))
    for n in (0, 1, 7, 64, 1000003):
        for w in (1, 2, 3, 8):
            r = [bench.shard_range(n, k, w) for k in range(w)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(r, r[1:]))
            assert max(b - a for a, b in r) - min(b - a for a, b in r) <= 1


class _HostShard:
    """Test stand-in for the per-rank GPU shard (numpy brute force + the merge rule of k_knn_merge_shards), so the
    world-size-2 host logic of distributed.ShardedKnnIndex -- index mapping, candidate gather layout -- runs on CPU."""

    def __init__(self, dim):
        self.pts = np.zeros((0, dim))

    def add(self, pts):
        self.pts = np.vstack([self.pts, pts])

    def topk(self, queries, k):
        nq = queries.shape[0]
        idx = np.full((nq, k), -1, np.int64)
        d2 = np.full((nq, k), np.inf)
        for i, q in enumerate(queries):
            d = ((self.pts - q) ** 2).sum(1)
            order = np.lexsort((np.arange(len(d)), d))[:k]
            idx[i, : len(order)] = order
            d2[i, : len(order)] = d[order]
        return torch.from_numpy(idx), torch.from_numpy(d2)

    def merge(self, cand_d2, cand_idx, world, nq, k):
        cd, ci = cand_d2.numpy(), cand_idx.numpy()
        out_i = np.full((nq, k), -1, np.int64)
        out_d = np.full((nq, k), np.inf)
        for q in range(nq):
            c = [(cd[g, q, j], ci[g, q, j] * world + g) for g in range(world) for j in range(k) if ci[g, q, j] >= 0]
            c.sort()
            for j, (d, gi) in enumerate(c[:k]):
                out_i[q, j], out_d[q, j] = gi, d
        return torch.from_numpy(out_i), torch.from_numpy(out_d)


def _knn_worker(rank, world, port, pts, queries, out_path):
    sys.path.insert(0, ROOT)
    import importlib

    import __graft_entry__ as g

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    g._load("sirrt_b200.distributed", os.path.join(g.PKG_DIR, "distributed.py"))
    D = sys.modules["sirrt_b200.distributed"]
    idx = D.ShardedKnnIndex(pts.shape[1], len(pts), _backend=_HostShard(pts.shape[1]))
    idx.add(pts[:1])          # batches of uneven size: the owner of a global index is index mod world
    idx.add(pts[1:10])
    idx.add(pts[10:])
    assert len(idx) == len(pts) and idx._b.pts.shape[0] == len(range(rank, len(pts), world))
    ti, td = idx.topk(queries, 5)
    ni, nd = idx.nearest(queries)
    rad = idx.radius(queries[0], 4.0)
    t = torch.arange(6 * 5, dtype=torch.int32).reshape(6, 5) + 100 * rank
    full = D.all_gather_interval_table(t)
    assert full.shape == (6 * world, 5) and all((full[6 * r: 6 * r + 6] == torch.arange(30, dtype=torch.int32).reshape(6, 5) + 100 * r).all() for r in range(world))
    if rank == 0:
        np.savez(out_path, ti=ti, td=td, ni=ni, nd=nd, rad=rad)
    dist.destroy_process_group()


def test_two_rank_sharded_knn(tmp_path):
    """kNN over a tree grown across two ranks == brute force over the whole set (ties -> lowest global index)."""
    rng = np.random.default_rng(11)
    pts = rng.uniform(-3.2, 3.2, (301, 6))
    pts[200] = pts[7]
    pts[150] = pts[7]    # duplicates on both shards
    queries = np.vstack([pts[7], rng.uniform(-3.2, 3.2, (20, 6))])
    out = str(tmp_path / "knn.npz")
    mp.spawn(_knn_worker, args=(2, 29533, pts, queries, out), nprocs=2, join=True)
    got = np.load(out)
    for i, q in enumerate(queries):
        d = ((pts - q) ** 2).sum(1)
        order = np.lexsort((np.arange(len(d)), d))
        assert np.array_equal(got["ti"][i], order[:5]) and np.array_equal(got["td"][i], d[order[:5]])
        assert got["ni"][i] == order[0] and got["nd"][i] == d[order[0]]
    d0 = ((pts - queries[0]) ** 2).sum(1)
    assert np.array_equal(got["rad"], np.nonzero(d0 < 4.0)[0]) and got["ni"][0] == 7
