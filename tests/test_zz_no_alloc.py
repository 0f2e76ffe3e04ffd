"""sirrt_knn_nearest_noalloc -- the 1-NN entry point that allocates nothing (scratch = a fixed file-scope device array; no
counterpart in the reference).  Collected LAST on purpose
(file name).

Checked through the C ABI against the existing entry point (sirrt_knn_nearest on a store object) on the same points: indices
and squared distances must be EQUAL (the tolerance of every kNN test here: same FP64 arithmetic, ties to the lowest index), on
two calls in a row -- the second runs on the scratch array the first one left behind -- plus the size limit.
torch only provides the device buffers of the test (the entry point takes device pointers and cannot stage host data
without allocating)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _call(pkg, torch, pts_dev, n, stride, dim, queries):
    q_dev = torch.from_numpy(np.ascontiguousarray(queries, np.float64)).cuda()
    nq = queries.shape[0]
    idx_dev = torch.full((max(nq, 1),), -7, dtype=torch.int64, device="cuda")
    d2_dev = torch.full((max(nq, 1),), -7.0, dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    stream = torch.cuda.current_stream().cuda_stream
    rc = pkg._lib.sirrt_knn_nearest_noalloc(0, stream, pts_dev.data_ptr(), n, stride, dim, q_dev.data_ptr(), nq, idx_dev.data_ptr(), d2_dev.data_ptr())
    torch.cuda.synchronize()
    return rc, idx_dev.cpu().numpy()[:nq], d2_dev.cpu().numpy()[:nq]


@pytest.mark.parametrize("dim,n", [(6, 20000), (7, 777), (3, 1)])
def test_noalloc_nearest_equals_store_nearest_twice(pkg, dim, n):
    torch = pytest.importorskip("torch")
    rng = np.random.default_rng(1234 + dim)
    pts = rng.uniform(-3.0, 3.0, (n, dim))
    if n > 100:
        pts[50] = pts[10]       # exact duplicates: the tie must go to the lower index
        pts[n - 1] = pts[n // 2]
    store = pkg.KnnIndex(dim, max(n, 16))
    store.add(pts)
    stride = n + 37             # a row stride larger than n, as a growable store has
    soa = np.zeros((dim, stride))
    soa[:, :n] = pts.T
    pts_dev = torch.from_numpy(soa).cuda()
    for call, nq in enumerate((256, 19)):   # two calls in a row; the second on the scratch the first left behind
        q = rng.uniform(-3.0, 3.0, (nq, dim))
        if n > 100:
            q[0] = pts[50]      # distance 0 to two points
            q[1] = pts[n - 1]
        want_idx, want_d2 = store.nearest(q)
        rc, idx, d2 = _call(pkg, torch, pts_dev, n, stride, dim, q)
        assert rc == 0, pkg._lib.sirrt_last_error()
        assert np.array_equal(idx, want_idx), (call, np.nonzero(idx != want_idx)[0][:8])
        assert np.array_equal(d2, want_d2), call
    if n > 100:
        assert idx[0] == 10 and idx[1] == n // 2
    store.close()


def test_noalloc_limits_and_empty(pkg):
    torch = pytest.importorskip("torch")
    dim = 6
    pts_dev = torch.zeros((dim, 64), dtype=torch.float64, device="cuda")
    q = np.zeros((257, dim))
    rc, _, _ = _call(pkg, torch, pts_dev, 64, 64, dim, q)   # more queries than the fixed scratch holds: rejected, nothing launched
    assert rc == -4 and b"SIRRT_KNN_NOALLOC_MAX_QUERIES" in pkg._lib.sirrt_last_error()
    rc, idx, d2 = _call(pkg, torch, pts_dev, 0, 64, dim, q[:3])  # empty point set: -1 / +inf, as sirrt_knn_nearest answers
    assert rc == 0 and (idx == -1).all() and np.isinf(d2).all()
    rc, _, _ = _call(pkg, torch, pts_dev, 65, 64, dim, q[:3])    # n beyond the row stride
    assert rc != 0
