"""world_size-2 gloo test of the N>1 path's host logic: contiguous shards of the configuration
batch, no data-path collective, results concatenated (all_gather of the fixed-width interval table
is the optional exchange step)."""
import os
import sys

import numpy as np
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
    n0, n1 = bench.shard_range(q.shape[1], rank, world)
    o = g.load_oracle().Oracle(pack)  # stands in for the per-rank GPU scene on this CPU-only box
    table, _ = o.safe_intervals_batch(np.ascontiguousarray(q[:, n0:n1]), 8)
    t = torch.from_numpy(table)
    parts = [torch.zeros_like(t) for _ in range(world)] if (n1 - n0) * world == q.shape[1] else None
    assert parts is not None
    dist.all_gather(parts, t)
    total = torch.tensor([n1 - n0], dtype=torch.int64)
    dist.all_reduce(total)
    if rank == 0:
        np.save(out_path, torch.cat(parts).numpy())
        assert int(total) == q.shape[1]
    dist.destroy_process_group()


def test_two_rank_shards_reassemble(tmp_path, sp, orc, packs):
    pk = packs("synA_8_50")
    q = sp.random_configs(sp.read_pack(pk).robot, 64, 3)
    out = str(tmp_path / "gathered.npy")
    mp.spawn(_worker, args=(2, 29531, pk, q, out), nprocs=2, join=True)
    full, _ = orc.Oracle(pk).safe_intervals_batch(q, 8)
    assert np.array_equal(np.load(out), full)


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
