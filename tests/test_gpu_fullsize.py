"""Full-size GPU tests: BASELINE.json configs[2] (7-DOF capsule arm, 2^20 configurations x 1000 frames x 64
moving boxes) through the C ABI.  The oracle cannot finish this size in seconds, so parity is established by
  (1) a bit-exact comparison with the oracle on a seeded sub-sample of the rows, and
  (2) size-independent properties of the full result: the interval table is exactly the complement-run
      decomposition of the bit rows (CollisionManager.cpp:507-546), results do not depend on how the batch is
      split into launches or on a row's position in the batch, the sweep kernel and the generic kernel agree,
      and padded frames beyond T never carry bits."""
import numpy as np
import pytest

from conftest import random_q

pytestmark = pytest.mark.gpu

N_FULL, T, C, KMAX = 1 << 20, 1000, 64, 8


@pytest.fixture(scope="module")
def full(pkg, sp, packs):
    pk = packs(f"synA_{C}_{T}")
    q = random_q(sp, pk, N_FULL, 42)
    dev = pkg.Scene(pk)
    table, bits, near = dev.safe_intervals_batch(q, KMAX)
    return pk, q, dev, table, bits, near


def _runs_from_bits(bits, t):
    """Vectorised complement runs of every row: (count, first lo, last hi) from the packed rows."""
    n, w = bits.shape
    b = np.unpackbits(bits.view(np.uint8).reshape(n, w * 4), axis=1, bitorder="little")[:, :t]
    safe = 1 - b.astype(np.int8)
    pad = np.zeros((n, 1), np.int8)
    d = np.diff(np.concatenate([pad, safe, pad], axis=1), axis=1)  # +1 at interval starts, -1 one past interval ends
    return (d == 1), (d == -1)


def test_full_size_table_is_complement_of_bits(full):
    pk, q, dev, table, bits, near = full
    assert bits.shape == (N_FULL, (T + 31) // 32) and table.shape == (N_FULL, 1 + 2 * KMAX)
    # padded frames never carry bits
    assert not (bits[:, -1] >> np.uint32(T - 32 * (bits.shape[1] - 1))).any()
    step = 1 << 16
    for n0 in range(0, N_FULL, step):
        starts, ends = _runs_from_bits(bits[n0:n0 + step], T)
        cnt = starts.sum(1)
        assert np.array_equal(cnt, table[n0:n0 + step, 0]), "interval count differs from the number of complement runs"
        assert np.array_equal(cnt, ends.sum(1))
        # k-th start / end for k < Kmax
        for k in range(KMAX):
            rows = np.nonzero(cnt > k)[0]
            if rows.size == 0:
                break
            cs = np.cumsum(starts[rows], axis=1)
            lo = (cs <= k).sum(1)  # index of the (k+1)-th start
            ce = np.cumsum(ends[rows], axis=1)
            hi = (ce <= k).sum(1) - 1  # d index of the (k+1)-th end marker is one past the last safe frame
            assert np.array_equal(lo, table[n0 + rows, 1 + 2 * k]) and np.array_equal(hi, table[n0 + rows, 2 + 2 * k])
        unused = np.arange(KMAX)[None, :] >= np.minimum(cnt, KMAX)[:, None]
        t2 = table[n0:n0 + step, 1:].reshape(-1, KMAX, 2)
        assert (t2[unused] == -1).all()
    dens = np.unpackbits(bits.view(np.uint8)).mean() * bits.shape[1] * 32 / T
    assert 0.05 < dens < 0.95, f"degenerate workload: colliding cell fraction {dens}"


def test_full_size_subsample_bit_exact_vs_oracle(full, orc):
    pk, q, dev, table, bits, near = full
    rng = np.random.default_rng(7)
    rows = np.sort(rng.choice(N_FULL, 768, replace=False))
    ref = orc.Oracle(pk)
    rtable, rbits = ref.safe_intervals_batch(np.ascontiguousarray(q[:, rows]), KMAX)
    assert np.array_equal(bits[rows], rbits), f"{int((bits[rows] != rbits).any(1).sum())} of 768 sampled rows differ"
    assert np.array_equal(table[rows], rtable)


def test_full_size_split_and_position_invariance(full, pkg):
    """Rows computed alone / in another order / in small launches equal the rows of the full batch."""
    pk, q, dev, table, bits, near = full
    rng = np.random.default_rng(11)
    perm = rng.permutation(N_FULL)[:50000]
    t2, b2, _ = dev.safe_intervals_batch(np.ascontiguousarray(q[:, perm]), KMAX)
    assert np.array_equal(b2, bits[perm]) and np.array_equal(t2, table[perm])
    dev.set_option(pkg.OPT_CHUNK_CONFIGS, 7777)  # odd launch size: partial warps / groups at every launch boundary
    t3, b3, _ = dev.safe_intervals_batch(np.ascontiguousarray(q[:, :100003]), KMAX)
    assert np.array_equal(b3, bits[:100003]) and np.array_equal(t3, table[:100003])


def test_full_size_generic_kernel_agrees(full, pkg):
    pk, q, dev, table, bits, near = full
    g = pkg.Scene(pk)
    g.set_option(pkg.OPT_USE_SWEEP, 0)
    sl = slice(500000, 500000 + 8192)
    t2, b2, _ = g.safe_intervals_batch(np.ascontiguousarray(q[:, sl]), KMAX)
    assert np.array_equal(b2, bits[sl]) and np.array_equal(t2, table[sl])


def test_full_size_near_contact_is_reported(full):
    pk, q, dev, table, bits, near = full
    st = dev.last_stats()
    assert near >= 0 and st["exact_tests"] > 0
    # the FP64 re-check is a thin band, not a second pass over everything
    assert st["rechecked"] < 0.01 * st["exact_tests"] + 1000


def test_pipelined_host_path_equals_plain_path_and_compact_table(full, pkg):
    """The 3-stream pipeline of the host-buffer path (H2D || kernels || D2H, blocks of 131 072 configurations) gives the
    plain single-stream path's table and bits at full size (the near-contact count within its scheduling tolerance); the int16 table is the int32 table narrowed."""
    pk, q, dev, table, bits, near = full   # computed with the pipeline on (the default for host buffers)
    plain = pkg.Scene(pk)
    plain.set_option(pkg.OPT_PIPELINE, 0)
    n = 400_003  # a ragged number of blocks
    qs = np.ascontiguousarray(q[:, :n])
    t0, b0, near0 = plain.safe_intervals_batch(qs, KMAX)
    assert np.array_equal(t0, table[:n]) and np.array_equal(b0, bits[:n])
    t1, b1, near1 = dev.safe_intervals_batch(qs, KMAX)
    assert np.array_equal(t1, t0) and np.array_equal(b1, b0)
    # near_contact_count counts the EVALUATED pair tests within 1e-6 m of contact.  A pair whose frame another warp lane has
    # already published as colliding is skipped, so the count depends (by a few per cent) on batching and scheduling; the
    # collision bits and tables never do.
    assert abs(near1 - near0) <= max(16, near0 // 10), (near1, near0)
    t1b, b1b, near1b = dev.safe_intervals_batch(qs, KMAX)
    assert np.array_equal(b1b, b1) and np.array_equal(t1b, t1), "the same call twice must give the same bits and table"
    assert abs(near1b - near1) <= max(16, near1 // 10)
    t2, _, _ = dev.safe_intervals_batch(qs, KMAX, want_bits=False)
    assert np.array_equal(t2, t0)
    t16, near16 = dev.safe_intervals_batch_i16(qs, KMAX)
    assert t16.dtype == np.int16 and np.array_equal(t16.astype(np.int32), t0) and abs(near16 - near1) <= max(16, near1 // 10)
    # one to two blocks' worth (a rank's share of the batch on 8 GPUs): cut in two halves so that the copies still overlap
    n2 = 140_001
    t3, b3, _ = dev.safe_intervals_batch(np.ascontiguousarray(q[:, :n2]), KMAX)
    assert np.array_equal(t3, table[:n2]) and np.array_equal(b3, bits[:n2])
    small16, _ = dev.safe_intervals_batch_i16(np.ascontiguousarray(q[:, :1000]), KMAX)   # a single block
    assert np.array_equal(small16.astype(np.int32), table[:1000])
