"""One process per GPU (`torch.distributed`, NCCL over NVLink): the two exchange steps the hot path has when a
SINGLE tree is grown across GPUs --

* `ShardedKnnIndex`: the vertex store of `MSIRRT/src/Tree.cpp:18-36` split over the ranks by insertion index
  (global index g lives on rank g mod G at local index g // G).  Every rank searches its own shard with the
  warp-cooperative brute force of `csrc/sirrt_knn.cuh`, the per-shard candidates (k of them per query) are
  all-gathered over NCCL and merged on every rank by `k_knn_merge_shards`: ascending (dist2, global index), so the
  answer is the one the single-GPU store (and nanoflann, `PlannerConnect.cpp:382-412`, up to ties) gives.
* `all_gather_interval_table`: the fixed-width safe-interval table `[N/G][1+2*Kmax]` of each rank's configuration
  shard gathered so that every rank holds all N rows.

PyTorch is plumbing here (device buffers + the NCCL collective); the searches and the merge are this
repository's CUDA kernels, called through the C ABI with device pointers.  There is no CPU path: the default
backend needs libsirrt_b200.so and a CUDA device.  `_backend` is a test seam for the world-size-2 gloo tests of
the host logic (index mapping, gather layout), never set by product code.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch
import torch.distributed as dist


def owner_of(global_index: int, world: int) -> tuple[int, int]:
    """(rank, local index) of a global vertex index."""
    return global_index % world, global_index // world


def local_slice(n0: int, n: int, rank: int, world: int) -> slice:
    """Rows of a batch of n points, appended at global index n0, that rank keeps (in order)."""
    first = (rank - n0) % world
    return slice(first, n, world)


class _CudaBackend:
    """Local shard on this rank's GPU + the device merge, through the C ABI."""

    def __init__(self, dim: int, capacity: int, device: int):
        import sys

        pkg = sys.modules[__package__] if __package__ else sys.modules["sirrt_b200"]
        self._pkg, self._lib, self.device = pkg, pkg._lib, device
        self.index = pkg.KnnIndex(dim, capacity, device)
        self.torch_device = torch.device("cuda", device)

    def add(self, pts: np.ndarray):
        self.index.add(pts)

    def topk(self, queries: np.ndarray, k: int):
        nq = queries.shape[0]
        idx = torch.empty((nq, k), dtype=torch.int64, device=self.torch_device)
        d2 = torch.empty((nq, k), dtype=torch.float64, device=self.torch_device)
        torch.cuda.current_stream(self.torch_device).synchronize()
        self._pkg._ck(self._lib.sirrt_knn_topk(self.index._h, queries.ctypes.data, nq, k, idx.data_ptr(), d2.data_ptr()))
        return idx, d2

    def merge(self, cand_d2: torch.Tensor, cand_idx: torch.Tensor, world: int, nq: int, k: int):
        out_i = torch.empty((nq, k), dtype=torch.int64, device=self.torch_device)
        out_d = torch.empty((nq, k), dtype=torch.float64, device=self.torch_device)
        st = torch.cuda.current_stream(self.torch_device).cuda_stream
        self._pkg._ck(self._lib.sirrt_knn_merge_shards(self.device, C.c_void_p(st), cand_d2.data_ptr(), cand_idx.data_ptr(), world, nq, k,
                                                       out_i.data_ptr(), out_d.data_ptr()))
        return out_i, out_d


class ShardedKnnIndex:
    """Exact FP64 kNN over a vertex set sharded across the ranks of `group` (all ranks call every method with the
    same arguments, as ranks growing one replicated tree do)."""

    def __init__(self, dim: int, capacity: int, device: int = 0, group=None, _backend=None):
        self.dim, self.group = dim, group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        if self.world > 64:
            raise ValueError("at most 64 shards (SIRRT_KNN_MAX_SHARDS)")
        self.n = 0
        self._b = _backend if _backend is not None else _CudaBackend(dim, (capacity + self.world - 1) // self.world + 1, device)

    def __len__(self):
        return self.n

    def add(self, pts):
        """Append points (same batch on every rank); global index = insertion order (Tree::add_vertex)."""
        pts = np.ascontiguousarray(pts, np.float64).reshape(-1, self.dim)
        mine = pts[local_slice(self.n, pts.shape[0], self.rank, self.world)]
        if mine.shape[0]:
            self._b.add(np.ascontiguousarray(mine))
        self.n += pts.shape[0]

    def topk(self, queries, k: int):
        """k nearest of each query over ALL shards: (global idx [nq][k], dist2 [nq][k]) ascending by (dist2, idx);
        -1 / inf past the number of stored points."""
        q = np.ascontiguousarray(queries, np.float64).reshape(-1, self.dim)
        nq = q.shape[0]
        li, ld = self._b.topk(q, k)
        cand_i = torch.empty((self.world * nq, k), dtype=torch.int64, device=li.device)   # [G][nq][k], rank-major
        cand_d = torch.empty((self.world * nq, k), dtype=torch.float64, device=ld.device)
        dist.all_gather_into_tensor(cand_i, li.contiguous(), group=self.group)   # G*k candidates per query
        dist.all_gather_into_tensor(cand_d, ld.contiguous(), group=self.group)
        gi, gd = self._b.merge(cand_d.view(self.world, nq, k), cand_i.view(self.world, nq, k), self.world, nq, k)
        return gi.cpu().numpy(), gd.cpu().numpy()

    def nearest(self, queries):
        """PlannerConnect::get_nearest_node (PlannerConnect.cpp:382-392) over the sharded store."""
        gi, gd = self.topk(queries, 1)
        return gi[:, 0], gd[:, 0]

    def radius(self, query, r2: float):
        """PlannerConnect::get_nearest_node_by_radius (:394-412): every global index with dist2 < r2, ascending."""
        q = np.ascontiguousarray(query, np.float64).reshape(1, self.dim)
        for k in (32, 128, 512, 1024):
            gi, gd = self.topk(q, k)
            inside = (gd[0] < r2) & (gi[0] >= 0)
            if not inside.all() or k >= self.n:
                return np.sort(gi[0][inside])
        raise ValueError("radius query with more than 1024 hits (the planner's r2 = 4 in a 6-D tree stays far below)")


def all_gather_interval_table(table_shard: torch.Tensor, group=None) -> torch.Tensor:
    """All ranks contribute the `[n_local][1+2*Kmax]` int32 interval table of their contiguous configuration shard
    (equal n_local on every rank) and receive the full `[N][1+2*Kmax]` table."""
    world = dist.get_world_size(group)
    out = torch.empty((world * table_shard.shape[0], table_shard.shape[1]), dtype=table_shard.dtype, device=table_shard.device)
    dist.all_gather_into_tensor(out, table_shard.contiguous(), group=group)
    return out
