"""(e) one tree grown across GPUs: the device merge of per-shard kNN candidates (single GPU, shards emulated by
several stores on one device) and the real thing over NCCL with one process per GPU (needs >= 2 GPUs:
`gpurun --gpus 2 -- python -m pytest tests/test_gpu_multi.py -m gpu`)."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


def _brute(pts, queries, k):
    out_i, out_d = [], []
    for q in queries:
        d = ((pts - q) ** 2).sum(1)
        order = np.lexsort((np.arange(len(d)), d))[:k]
        out_i.append(order)
        out_d.append(d[order])
    return np.array(out_i), np.array(out_d)


@pytest.mark.parametrize("G,k", [(2, 1), (3, 5), (8, 32)])
def test_merge_shards_kernel(pkg, G, k):
    import torch

    rng = np.random.default_rng(3)
    n, dim, nq = 4001, 6, 97
    pts = rng.uniform(-3.2, 3.2, (n, dim))
    pts[1000] = pts[3]
    pts[2001] = pts[3]
    queries = np.vstack([pts[3], rng.uniform(-3.2, 3.2, (nq - 1, dim))])
    shards = [pkg.KnnIndex(dim, n // G + 2) for _ in range(G)]
    cand_i = torch.empty((G, nq, k), dtype=torch.int64, device="cuda")
    cand_d = torch.empty((G, nq, k), dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    for g, s in enumerate(shards):
        s.add(np.ascontiguousarray(pts[g::G]))
        pkg._ck(pkg._lib.sirrt_knn_topk(s._h, queries.ctypes.data, nq, k, cand_i[g].data_ptr(), cand_d[g].data_ptr()))
    out_i = torch.empty((nq, k), dtype=torch.int64, device="cuda")
    out_d = torch.empty((nq, k), dtype=torch.float64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    pkg._ck(pkg._lib.sirrt_knn_merge_shards(0, C.c_void_p(st), cand_d.data_ptr(), cand_i.data_ptr(), G, nq, k, out_i.data_ptr(), out_d.data_ptr()))
    torch.cuda.synchronize()
    bi, bd = _brute(pts, queries, k)
    assert np.array_equal(out_i.cpu().numpy(), bi) and np.array_equal(out_d.cpu().numpy(), bd)
    # and it equals the single-store answer
    one = pkg.KnnIndex(dim, n)
    one.add(pts)
    si, sd = one.topk(queries, k)
    assert np.array_equal(si, bi) and np.array_equal(sd, bd)


def test_merge_shards_short_lists(pkg):
    """fewer stored points than k: the tail is -1 / inf, as the single store reports it"""
    import torch

    pts = np.random.default_rng(4).uniform(-1, 1, (5, 6))
    G, k = 4, 8
    shards = [pkg.KnnIndex(6, 8) for _ in range(G)]
    cand_i = torch.empty((G, 1, k), dtype=torch.int64, device="cuda")
    cand_d = torch.empty((G, 1, k), dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    q = np.zeros((1, 6))
    for g, s in enumerate(shards):
        if len(pts[g::G]):
            s.add(np.ascontiguousarray(pts[g::G]))
        pkg._ck(pkg._lib.sirrt_knn_topk(s._h, q.ctypes.data, 1, k, cand_i[g].data_ptr(), cand_d[g].data_ptr()))
    out_i = torch.empty((1, k), dtype=torch.int64, device="cuda")
    out_d = torch.empty((1, k), dtype=torch.float64, device="cuda")
    pkg._ck(pkg._lib.sirrt_knn_merge_shards(0, C.c_void_p(torch.cuda.current_stream().cuda_stream), cand_d.data_ptr(), cand_i.data_ptr(), G, 1, k,
                                            out_i.data_ptr(), out_d.data_ptr()))
    torch.cuda.synchronize()
    bi, bd = _brute(pts, q, 5)
    oi, od = out_i.cpu().numpy()[0], out_d.cpu().numpy()[0]
    assert np.array_equal(oi[:5], bi[0]) and np.array_equal(od[:5], bd[0]) and (oi[5:] == -1).all() and np.isinf(od[5:]).all()


def _nccl_worker(rank, world, port, pack, pts, queries, q_cfg, out_path):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist

    import __graft_entry__ as g

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    pkg = g.load_package()
    D = g._load("sirrt_b200.distributed", os.path.join(g.PKG_DIR, "distributed.py"))
    idx = D.ShardedKnnIndex(pts.shape[1], len(pts), device=rank)
    idx.add(pts[:1])
    idx.add(pts[1:1000])
    for p in pts[1000:1010]:
        idx.add(p[None])
    idx.add(pts[1010:])
    ti, td = idx.topk(queries, 7)
    ni, nd = idx.nearest(queries)
    rad = idx.radius(queries[0], 4.0)
    # interval tables of contiguous configuration shards, all-gathered so that every rank holds every row
    N = q_cfg.shape[1]
    n0, n1 = rank * (N // world), (rank + 1) * (N // world)
    scene = pkg.Scene(pack, rank)
    table, _, _ = scene.safe_intervals_batch(np.ascontiguousarray(q_cfg[:, n0:n1]), 8, want_bits=False)
    full = D.all_gather_interval_table(torch.from_numpy(table).cuda())
    torch.cuda.synchronize()
    np.savez(out_path % rank, ti=ti, td=td, ni=ni, nd=nd, rad=rad, table=full.cpu().numpy())
    dist.destroy_process_group()


def test_sharded_knn_and_table_gather_over_nccl(pkg, orc, sp, packs, tmp_path):
    import torch
    import torch.multiprocessing as mp

    world = min(torch.cuda.device_count(), 8)
    if world < 2:
        pytest.skip("needs >= 2 GPUs")
    rng = np.random.default_rng(17)
    pts = rng.uniform(-3.2, 3.2, (30011, 6))
    pts[20000] = pts[5]
    pts[1003] = pts[5]
    queries = np.vstack([pts[5], rng.uniform(-3.2, 3.2, (64, 6))])
    pk = packs("synA_32_200")
    q_cfg = sp.random_configs(sp.read_pack(pk).robot, 512 * world, 23)
    out = str(tmp_path / "r%d.npz")
    mp.spawn(_nccl_worker, args=(world, 29541, pk, pts, queries, q_cfg, out), nprocs=world, join=True)
    bi, bd = _brute(pts, queries, 7)
    d0 = ((pts - queries[0]) ** 2).sum(1)
    rtable, _ = orc.Oracle(pk).safe_intervals_batch(q_cfg, 8)
    for r in range(world):
        got = np.load(out % r)
        assert np.array_equal(got["ti"], bi) and np.array_equal(got["td"], bd)
        assert np.array_equal(got["ni"], bi[:, 0]) and np.array_equal(got["nd"], bd[:, 0])
        assert np.array_equal(got["rad"], np.nonzero(d0 < 4.0)[0])
        assert np.array_equal(got["table"], rtable)


@pytest.mark.parametrize("name,n,devices", [("synA_32_200", 20011, [0, 0]), ("synA_32_200", 10007, [0, 0, 0]), ("synB_16_100", 5, [0, 0, 0]),
                                            ("flying_boxes", 101, [0, 0, 0]), ("multiple_robots_cyl", 333, [0, 0])])
def test_multi_scene_splits_one_batch_like_a_single_scene(pkg, sp, packs, name, n, devices):
    """sirrt_multi_create / sirrt_multi_safe_intervals_batch (the in-process form of the §8e partition): one batch cut into
    G contiguous blocks, one scene replica and host thread per block -- G = 2 and 3, N not divisible by G, N < G blocks
    with empty shards, sweep and generic kernels.  Replicas on the SAME device exercise the slicing, the table / bits
    offsets and the summed near-contact count without needing a multi-GPU box; with more GPUs the replicas spread out."""
    import torch

    nd = torch.cuda.device_count()
    devs = [d % max(1, nd) for d in range(len(devices))] if nd >= len(devices) else devices
    pk = packs(name)
    q = sp.random_configs(sp.read_pack(pk).robot, n, 61)
    one = pkg.Scene(pk, 0)
    table, bits, near = one.safe_intervals_batch(q, 8)
    multi = pkg.MultiScene(pk, devs)
    mtable, mbits, mnear = multi.safe_intervals_batch(q, one.words, 8)
    assert np.array_equal(mtable, table)
    assert np.array_equal(mbits, bits)
    assert near // 3 - 8 <= mnear <= 3 * near + 8  # evaluated near-contact pair tests: depends slightly on the batching
    mtable2, none_bits, _ = multi.safe_intervals_batch(q, one.words, 8, want_bits=False)
    assert none_bits is None and np.array_equal(mtable2, table)
