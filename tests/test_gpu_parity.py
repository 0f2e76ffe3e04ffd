"""GPU parity tests: the CUDA path, called through the C ABI (libsirrt_b200.so), against the CPU
oracle on identical inputs.  Bar: collision bits and interval end points BIT-EXACT (the FP32 fast
path re-evaluates every pair inside its uncertainty band in FP64 with the oracle's formulas, so no
tolerance is needed; pair tests within 1e-6 m of contact are counted and reported)."""
import numpy as np
import pytest

from conftest import random_q

pytestmark = pytest.mark.gpu


def _compare(pkg, orc, pack, q, kmax=8, mode=0):
    dev = pkg.Scene(pack, 0, mode)
    dev.set_option(pkg.OPT_SWEEP_MIN_CONFIGS, 1)  # small batches too go through k_sweep where it applies
    ref = orc.Oracle(pack, mode)
    table, bits, near = dev.safe_intervals_batch(q, kmax)
    rtable, rbits = ref.safe_intervals_batch(q, kmax)
    bad = np.nonzero((bits != rbits).any(1))[0]
    assert bad.size == 0, f"{bad.size} rows differ, first {bad[:5]}; stats {dev.last_stats()}"
    assert np.array_equal(table, rtable)
    return dev, ref, bits, near


@pytest.mark.parametrize("name,n", [("synA_32_200", 512), ("synB_32_200", 512), ("synC_32_200", 512), ("synA_64_1000", 96),
                                    ("synA_5_33", 64), ("synB_3_1", 64), ("synA_70_64", 128), ("synA_32_200", 20011),
                                    ("synB_16_100", 9000), ("synC_16_100", 9000)])
def test_synthetic_sweeps_bit_exact(pkg, orc, sp, packs, name, n):
    pk = packs(name)
    dev, ref, bits, near = _compare(pkg, orc, pk, random_q(sp, pk, n, 13))
    assert bits.any() and not bits.all()


def test_generic_kernel_equals_sweep_kernel(pkg, orc, sp, packs):
    pk = packs("synB_32_200")
    q = random_q(sp, pk, 256, 21)
    a = pkg.Scene(pk)
    a.set_option(pkg.OPT_SWEEP_MIN_CONFIGS, 1)
    b = pkg.Scene(pk)
    b.set_option(pkg.OPT_USE_SWEEP, 0)
    ta, ba, _ = a.safe_intervals_batch(q, 8)
    tb, bb, _ = b.safe_intervals_batch(q, 8)
    assert np.array_equal(ba, bb) and np.array_equal(ta, tb)


@pytest.mark.parametrize("name,n", [("mass_test_cyl", 2048), ("mass_test", 2048), ("flying_boxes", 2048), ("scene_task", 16)])
def test_reference_scenes_bit_exact(pkg, orc, sp, packs, name, n):
    """>= 2 000 configurations per hull scene (1.2e6 - 3.7e6 cells each): the FP32 GJK + FP64 certificate + support-cell
    path is the riskiest code of the library"""
    pk = packs(name)
    _compare(pkg, orc, pk, random_q(sp, pk, n, 17))


@pytest.mark.parametrize("name,n", [("multiple_robots_cyl", 2048), ("multiple_robots", 2048)])
@pytest.mark.parametrize("mode", [0, 1])
def test_robot_obstacles_both_parity_modes(pkg, orc, sp, packs, name, n, mode):
    """QUIRK 7: REFERENCE_EXACT reproduces the functor mis-indexing, FIXED tests true geometry."""
    pk = packs(name)
    q = random_q(sp, pk, n, 19)
    dev, ref, bits, _ = _compare(pkg, orc, pk, q, mode=mode)
    # per-frame semantics (CollisionCallback: shifted functor only in REFERENCE_EXACT)
    qd = q.T.astype(np.float64)
    fb = dev.collision_bits_f64(qd[:8], pkg.SEM_FRAME)
    for i in range(8):
        rb, _ = ref.collision_bits(qd[i], orc.SEM_FRAME)
        assert np.array_equal(fb[i], rb)


def test_golden_vectors_through_the_c_abi(pkg, packs, golden_vectors):
    for name, ent in golden_vectors.items():
        dev = pkg.Scene(packs(name))
        for q, iv, fc in zip(ent["configs"], ent["intervals"], ent["frame_checks"]):
            assert dev.safe_intervals(q) == [tuple(x) for x in iv], name
            for f, c in fc:
                assert int(dev.check_frame(q, f)) == c


def test_single_call_dropins(pkg, orc, sp, packs):
    pk = packs("synB_32_200")
    dev, ref = pkg.Scene(pk), orc.Oracle(pk)
    q = random_q(sp, pk, 24, 23).T.astype(np.float64)
    for qq in q:
        assert dev.safe_intervals(qq) == ref.safe_intervals(qq)
        for f in (0, 57, 199):
            assert dev.check_frame(qq, f) == ref.check_frame(qq, f)
        assert dev.check_frame(qq, 200) and dev.check_frame(qq, -1)  # out of range answers "collided"
        R, t = dev.fk(qq)
        R0, t0 = ref.fk(qq)
        assert np.abs(R - R0).max() < 1e-14 and np.abs(t - t0).max() < 1e-14
    # capacity overflow: true count returned with the error code
    many = [qq for qq in q if len(ref.safe_intervals(qq)) > 1]
    if many:
        with pytest.raises(pkg.SirrtError) as e:
            dev.safe_intervals(many[0], cap=1)
        assert e.value.code == -5


def test_check_frames_batch(pkg, orc, sp, packs):
    for name in ("synB_32_200", "flying_boxes", "multiple_robots_cyl"):
        pk = packs(name)
        dev, ref = pkg.Scene(pk), orc.Oracle(pk)
        n = 600 if name.startswith("syn") else 96
        q = random_q(sp, pk, n, 29)
        frames = np.random.default_rng(3).integers(-2, ref.n_frames + 2, n).astype(np.int32)
        got = dev.check_frames_batch(q, frames)
        exp = ref.check_frames_batch(q, frames)
        assert np.array_equal(got, exp), name
        got64 = dev.check_frames_batch_f64(q.T.astype(np.float64), frames)
        assert np.array_equal(got64, exp)


@pytest.mark.parametrize("name,mode", [("synA_32_200", 0), ("flying_boxes", 0), ("mass_test", 0), ("multiple_robots", 0), ("multiple_robots", 1)])
def test_edge_check_batch(pkg, orc, sp, packs, name, mode):
    """PlannerConnect::is_collision_motion (PlannerConnect.cpp:862-897) through sirrt_edge_check_batch on the capsule
    scene AND on the hull / robot-obstacle scenes (per-frame semantics: the QUIRK-7 functor in REFERENCE_EXACT)"""
    pk = packs(name)
    dev, ref = pkg.Scene(pk, 0, mode), orc.Oracle(pk, mode)
    rng = np.random.default_rng(31)
    n = 200
    T = ref.n_frames
    q0 = random_q(sp, pk, n, 37).T.astype(np.float64)
    q1 = q0 + rng.normal(0, 0.4, q0.shape)
    t0 = rng.uniform(0, 0.75 * T, n)
    t1 = t0 + rng.uniform(-5, 60, n)
    first, coord = dev.edge_check_batch(q0, q1, t0, t1, vmax=3.1415)
    kinds = set()
    for e in range(n):
        s, cc = ref.edge_check(q0[e], q1[e], t0[e], t1[e], 3.1415)
        assert first[e] == s, (e, first[e], s)
        if s >= 0:
            assert np.array_equal(coord[e], cc)
        kinds.add(min(s, 0) if s < 0 else 1)
    assert {-3, -2, -1} <= kinds and (1 in kinds or name == "mass_test"), kinds


def test_get_distances(pkg, orc, sp, packs):
    for name in ("synB_32_200", "flying_boxes"):
        pk = packs(name)
        dev, ref = pkg.Scene(pk), orc.Oracle(pk)
        q = random_q(sp, pk, 12, 41).T.astype(np.float64)
        for i, qq in enumerate(q):
            f = (37 * i) % ref.n_frames
            col, d = dev.get_distances(qq, f)
            rcol, rd = ref.get_distances(qq, f)
            assert col == bool(rcol)
            k = np.isfinite(d) & (d < 1e299)
            assert np.allclose(d[k], rd[k], atol=1e-9), (d, rd)


@pytest.mark.parametrize("name", ["synB_32_200", "synC_32_200", "flying_boxes", "mass_test"])
def test_get_distances_nearest_points(pkg, orc, sp, packs, name):
    """a9: sirrt_get_distances_ex -- per (link, obstacle) the signed distance and the closest point pair (what
    DistanceCallback collects, CollisionManager.cpp:568-598).  Against the C oracle: pair distances.  Against the
    independent oracle2 (projection / scalar minimisation / NNLS-QP witness points): the points span the distance, each
    lies on its own shape, and where the closest pair is unique they coincide."""
    import os
    import __graft_entry__ as g

    o2 = g._load("sirrt_oracle2", os.path.join(g.ROOT, "oracle", "oracle2.py"))
    pk = packs(name)
    dev, ref, O2 = pkg.Scene(pk), orc.Oracle(pk), o2.Oracle2(pk)
    q = random_q(sp, pk, 6, 43).T.astype(np.float64)
    n_unique = 0
    for i, qq in enumerate(q):
        f = (53 * i + 7) % ref.n_frames
        col, d, pd, pl, po = dev.get_distances_ex(qq, f)
        col0, d0 = dev.get_distances(qq, f)
        assert col == col0 and np.array_equal(d, d0)
        links = O2.links(qq)
        for l in range(dev.n_links):
            for c in range(O2.C):
                dref = ref.pair_distance(qq, l, c, f)
                if dref > 1e-6:  # separated pairs: penetrating ones have no meaningful witness (the reference returns before reading them)
                    assert abs(pd[l, c] - dref) < 1e-9, (name, l, c, pd[l, c], dref)
                    if dref > 0.5:
                        continue  # far pairs add nothing; keep the solver count down
                    ob = O2.obstacle(c, f)
                    # the packs carry FP32 obstacle rotations (orthonormal to ~1e-7): a distance evaluated in the obstacle's
                    # frame and the world-frame span of the two points differ by that much
                    assert abs(np.linalg.norm(pl[l, c] - po[l, c]) - dref) < 3e-7 * max(1.0, dref)
                    pt = lambda x: o2.Posed(o2.Geom(o2.SPHERE, radius=0.0), np.block([[np.eye(3), x[:, None]], [np.zeros((1, 3)), np.ones((1, 1))]]))
                    assert abs(o2.signed_distance(pt(pl[l, c]), links[l])) < 1e-7, "link point is not on the link surface"
                    assert abs(o2.signed_distance(pt(po[l, c]), ob)) < 1e-7, "obstacle point is not on the obstacle surface"
                    d2, a2, b2 = o2.closest_points(links[l], ob)
                    assert abs(d2 - dref) < 1e-7
                    if np.linalg.norm(a2 - pl[l, c]) < 1e-6 and np.linalg.norm(b2 - po[l, c]) < 1e-6:
                        n_unique += 1
    assert n_unique > 0, "no pair whose witness points agree with the second oracle"


def test_edge_cases(pkg, orc, sp, packs):
    pk = packs("synA_32_200")
    dev, ref = pkg.Scene(pk), orc.Oracle(pk)
    # empty batch
    t, b, near = dev.safe_intervals_batch(np.zeros((7, 0), np.float32), 4)
    assert t.shape == (0, 9) and b.shape == (0, 7) and near == 0
    # Kmax smaller than the interval count: count is the true count, only Kmax pairs stored
    q = random_q(sp, pk, 128, 43)
    t1, _, _ = dev.safe_intervals_batch(q, 1)
    t8, _ = ref.safe_intervals_batch(q, 8)
    assert np.array_equal(t1[:, 0], t8[:, 0]) and (t1[:, 0] > 1).any()
    assert np.array_equal(t1[t8[:, 0] > 0, 1:3], t8[t8[:, 0] > 0, 1:3])
    # a tiny re-check work list forces the overflow -> split -> redo path; results must not change
    dev2 = pkg.Scene(pk)
    dev2.set_band(5e-3)
    dev2.set_option(pkg.OPT_WORKLIST_CAP, 64)
    tt, bb, _ = dev2.safe_intervals_batch(q, 8)
    assert np.array_equal(tt, t8)
    # a wide band sends everything near contact to FP64: same answer, more rechecks
    assert dev2.last_stats()["rechecked"] > 0
    # device-resident inputs / outputs (torch tensors) through the same entry point
    import torch

    qd = torch.from_numpy(q).cuda()
    table = torch.zeros((q.shape[1], 17), dtype=torch.int32, device="cuda")
    dev.set_stream(torch.cuda.current_stream().cuda_stream)
    dev.safe_intervals_batch_raw(qd.data_ptr(), q.shape[1], table.data_ptr(), 8, None)
    assert np.array_equal(table.cpu().numpy(), t8)


def test_knn(pkg, orc):
    rng = np.random.default_rng(5)
    pts = rng.uniform(-3.2, 3.2, (20000, 6))
    pts[1234] = pts[77]   # exact duplicates: the lower index must win
    pts[15000] = pts[77]
    idx = pkg.KnnIndex(6, 30000)
    idx.add(pts[:5000])
    idx.add(pts[5000:])   # incremental insertion keeps insertion order = index (Tree::add_vertex)
    assert len(idx) == 20000
    qs = np.vstack([pts[77], pts[77] + 1e-9, rng.uniform(-3.2, 3.2, (300, 6))])
    gi, gd = idx.nearest(qs)
    ri, rd = orc.knn_nearest(pts, qs)
    assert np.array_equal(gi, ri) and np.array_equal(gd, rd)  # FP64, same accumulation order: bit-exact
    assert gi[0] == 77
    for k in (0, 5, 50):
        got, n = idx.radius(qs[k], 4.0)
        exp = orc.knn_radius(pts, qs[k], 4.0)
        assert n == len(exp) and np.array_equal(got, exp)
    ti, td = idx.topk(qs[:40], 8)
    d2 = ((pts[None] - qs[:40, None]) ** 2).sum(-1)
    for k in range(40):
        order = np.lexsort((np.arange(len(pts)), d2[k]))[:8]
        assert np.array_equal(np.sort(d2[k][ti[k]]), np.sort(d2[k][order]))
        assert np.array_equal(ti[k, :1], gi[k : k + 1])
    # empty index
    e = pkg.KnnIndex(6, 10)
    i0, _ = e.nearest(qs[:2])
    assert (i0 == -1).all() and e.radius(qs[0], 1.0)[1] == 0


def _topk_ref(pts, q, k):
    """ascending (distance, index), distance accumulated in dimension order without FMA (float64 numpy does exactly that)"""
    d2 = np.zeros(len(pts))
    for j in range(pts.shape[1]):
        e = q[j] - pts[:, j]
        d2 = d2 + e * e
    order = np.lexsort((np.arange(len(pts)), d2))[:k]
    return order, d2[order]


@pytest.mark.parametrize("dim", [1, 2, 3, 6, 7, 8, 9, 16])
def test_knn_tiled_multi_query_is_exact(pkg, orc, dim):
    """The throughput path (k_knn_nearest_tiled: FP32 filter with a rigorous error bound + FP64 verification) returns the
    oracle's index and distance bit for bit -- random data, exact duplicates, clusters whose members differ far below FP32
    resolution, large offsets -- for store sizes around the tile size and query counts around the warp size; the store grows
    from a tiny initial capacity.  PlannerConnect.cpp:382-392 (ties -> lowest index)."""
    rng = np.random.default_rng(100 + dim)
    for n, kind in [(1, "u"), (255, "u"), (256, "dup"), (257, "cluster"), (5000, "u"), (5003, "cluster"), (3000, "offset"), (4100, "dup")]:
        pts = rng.uniform(-3.2, 3.2, (n, dim))
        if kind == "dup" and n > 8:
            pts[n // 2:] = pts[: n - n // 2]              # every point twice: the lower index must win
        elif kind == "cluster":
            c = rng.uniform(-3, 3, (8, dim))
            pts = c[rng.integers(0, 8, n)] + 1e-9 * rng.standard_normal((n, dim))   # FP32 cannot tell the members apart
        elif kind == "offset":
            pts = pts * 1e-3 + 1000.0
        idx = pkg.KnnIndex(dim, 4)
        for a in range(0, n, 1777):
            idx.add(pts[a : a + 1777])
        assert len(idx) == n
        for nq in (2, 31, 33, 300):
            qs = rng.uniform(-3.2, 3.2, (nq, dim)) if kind != "offset" else 1000.0 + 1e-3 * rng.uniform(-3.2, 3.2, (nq, dim))
            qs[0] = pts[rng.integers(0, n)]               # a query on a stored point: distance 0, duplicates tie
            if kind == "cluster":
                qs[1:] = pts[rng.integers(0, n, nq - 1)] + 1e-9 * rng.standard_normal((nq - 1, dim))
            gi, gd = idx.nearest(qs)
            ri, rd = orc.knn_nearest(pts, qs)
            assert np.array_equal(gi, ri), (dim, n, kind, nq, np.flatnonzero(gi != ri)[:5])
            assert np.array_equal(gd, rd)
        idx.close()


def test_knn_single_pass_topk_is_exact(pkg):
    """k <= 32: per-warp sorted lists merged across warps and splits give the lexicographic (distance, index) order exactly
    -- also with duplicates, with fewer points than k (missing entries: index -1, distance inf), and for k > 32 (the k-pass
    kernel)."""
    rng = np.random.default_rng(9)
    for n, dim in [(5, 6), (31, 3), (1000, 6), (20011, 6), (70000, 7)]:
        pts = rng.uniform(-3.2, 3.2, (n, dim))
        if n > 100:
            pts[n // 3 : n // 3 + 50] = pts[:50]
        idx = pkg.KnnIndex(dim, 16)
        idx.add(pts)
        qs = np.vstack([pts[:3], rng.uniform(-3.2, 3.2, (21, dim))])
        for k in (1, 8, 32, 40):
            ti, td = idx.topk(qs, k)
            for j in range(len(qs)):
                o, d = _topk_ref(pts, qs[j], k)
                m = len(o)
                assert np.array_equal(ti[j, :m], o), (n, dim, k, j)
                assert np.array_equal(td[j, :m], d)
                assert (ti[j, m:] == -1).all() and np.isinf(td[j, m:]).all()
        idx.close()


def test_collision_manager_mirror(pkg, orc, packs, sp):
    """The reference-facing operator interface (MDP::CollisionManager method names)."""
    pk = packs("flying_boxes")
    cm = pkg.CollisionManager(pk)
    ref = orc.Oracle(pk)
    sc = sp.read_pack(pk)
    lim = cm.get_planned_robot_limits()
    assert len(lim) == 6 and abs(lim[1][0] - (-2.059)) < 1e-6
    assert cm.get_safe_intervals(sc.start_configuration) == ref.safe_intervals(sc.start_configuration)
    assert cm.check_collision_frame(sc.start_configuration, 0) is False          # check_scene.cpp:74
    assert cm.check_collision(sc.end_configuration, (sc.n_frames - 1) / sc.fps + 1e-3) is False  # :83
    assert cm.check_robot_selfcollision(sc.start_configuration) is False
    low = cm.get_goal_frame_low_bound()
    col = [ref.check_frame(sc.end_configuration, f) for f in range(sc.n_frames)]
    assert low == (max(i for i, c in enumerate(col) if c) + 1 if any(col) else 0)


@pytest.mark.parametrize("key", ["strrt_mass_test", "strrt_flying_boxes", "drgbt_mass_test", "rgbtstar_mass_test", "drgbt_flying_boxes",
                                 "drgbt_multiple_robots", "strrt_multiple_robots"])
def test_reference_logged_trajectories_free_on_gpu(pkg, orc, packs, logged_cells, key):
    """Sanity check (not a pin): cells of trajectories logged by the project's older pipeline (different checker, see
    tests/golden/make_logged_paths.py) must be FREE through the C ABI too, and equal to the oracle cell by cell (robot scene: FIXED mode, see
    tests/test_oracle.py::test_reference_logged_trajectories_are_free)."""
    scene, q, frames = logged_cells(key)
    mode = pkg.PARITY_FIXED if scene == "multiple_robots" else pkg.PARITY_REFERENCE_EXACT
    dev = pkg.Scene(packs(scene), 0, mode)
    got = np.asarray(dev.check_frames_batch_f64(q, frames)).astype(bool)
    assert got.shape[0] == q.shape[0] and not got.any(), (key, int(got.sum()))
    ref = orc.Oracle(packs(scene), mode)
    step = max(1, len(q) // 200)
    assert [bool(ref.check_frame(qq, int(f))) for qq, f in zip(q[::step], frames[::step])] == got[::step].tolist()


@pytest.mark.parametrize("name", ["flying_boxes", "mass_test", "multiple_robots"])
def test_deferred_hull_tasks_equal_in_warp_gjk(pkg, orc, sp, packs, name):
    """Hull scenes: small batches hand their GJK pair tests to the machine-wide task kernel (k_gjk_tasks); the
    rows, the cells and the single-call intervals must equal the in-warp path and the oracle bit for bit."""
    pk = packs(name)
    q = random_q(sp, pk, 12, 29)
    a, b = pkg.Scene(pk), pkg.Scene(pk)
    b.set_option(pkg.OPT_DEFER_MAX_CONFIGS, 0)
    ta, ba, _ = a.safe_intervals_batch(q, 8)
    tb, bb, _ = b.safe_intervals_batch(q, 8)
    assert np.array_equal(ba, bb) and np.array_equal(ta, tb)
    ref = orc.Oracle(pk)
    qd = q.T.astype(np.float64).copy()
    fr = (np.arange(12, dtype=np.int32) * 131) % ref.n_frames
    ca, cb = a.check_frames_batch_f64(qd, fr), b.check_frames_batch_f64(qd, fr)
    assert np.array_equal(ca, cb)
    assert [bool(x) for x in ca] == [bool(ref.check_frame(qq, int(f))) for qq, f in zip(qd, fr)]
    for qq in qd[:4]:
        assert a.safe_intervals(qq) == ref.safe_intervals(qq) == b.safe_intervals(qq)
    fa, fb = a.collision_bits_f64(qd, pkg.SEM_FRAME), b.collision_bits_f64(qd, pkg.SEM_FRAME)
    assert np.array_equal(fa, fb)


def test_knn_single_query_latency_path(pkg, orc):
    """The planner's pattern (Tree::add_vertex then one nearest / one radius query per iteration): single-launch
    kernels, answers through the pinned block, asynchronous insertion ring (wraps every 64 points)."""
    rng = np.random.default_rng(11)
    pts = rng.uniform(-3.2, 3.2, (700, 6))
    pts[300] = pts[5]
    idx = pkg.KnnIndex(6, 4096)
    for n in range(len(pts)):
        idx.add(pts[n][None, :])
        if n % 7 == 0 or n > 690:
            q = rng.uniform(-3.2, 3.2, (1, 6)) if n % 2 else pts[5][None, :] + 1e-12
            gi, gd = idx.nearest(q)
            ri, rd = orc.knn_nearest(pts[: n + 1], q)
            assert gi[0] == ri[0] and gd[0] == rd[0], n
            got, cnt = idx.radius(q[0], 9.0)
            exp = orc.knn_radius(pts[: n + 1], q[0], 9.0)
            assert cnt == len(exp) and np.array_equal(got, exp), n
    big = rng.uniform(-1, 1, (3000, 6))  # more hits than the pinned block's first slice: second copy path
    idx2 = pkg.KnnIndex(6, 4096)
    idx2.add(big)
    got, cnt = idx2.radius(np.zeros(6), 100.0)
    assert cnt == 3000 and np.array_equal(got, np.arange(3000))
    got, cnt = idx2.radius(np.zeros(6), 0.5)
    assert np.array_equal(got, orc.knn_radius(big, np.zeros(6), 0.5))


def _random_motions(sp, pk, n, seed, fps_frames):
    rng = np.random.default_rng(seed)
    q0 = random_q(sp, pk, n, seed + 1).T.astype(np.float64)
    q1 = q0 + rng.normal(0, 0.25, q0.shape)
    horizon = fps_frames[1] / fps_frames[0]
    t0 = rng.uniform(0, 0.8 * horizon, n)
    t1 = t0 + rng.uniform(-0.2, 1.5, n)
    return q0, q1, t0, t1


@pytest.mark.parametrize("name", ["synA_32_200", "synB_32_200", "flying_boxes", "mass_test"])
@pytest.mark.parametrize("variant", [0, 1])
def test_motion_check_batch_matches_oracle(pkg, orc, sp, packs, name, variant):
    """f4: SpaceTimeMotionValidator::checkMotion (both overloads) for a batch of motions in one call vs the oracle's
    restatement of SpaceTimeMotionValidator.cpp:22-127: result code and lastValid fraction identical."""
    pk = packs(name)
    sc = sp.read_pack(pk)
    dev, ref = pkg.Scene(pk), orc.Oracle(pk)
    n = 160 if name.startswith("syn") else 48
    q0, q1, t0, t1 = _random_motions(sp, pk, n, 91, (sc.fps, sc.n_frames))
    res, frac = dev.motion_check_batch(q0, q1, t0, t1, 3.1415, variant)
    kinds = set()
    for e in range(n):
        r, f = ref.motion_check(q0[e], q1[e], t0[e], t1[e], 3.1415, variant)
        assert res[e] == r and frac[e] == f, (name, variant, e, res[e], r, frac[e], f)
        kinds.add(min(r, 0) if r < 0 else 1)
    assert {-3, -2, -1} <= kinds, kinds
    # empty batch
    r0, f0 = dev.motion_check_batch(np.zeros((0, dev.dof)), np.zeros((0, dev.dof)), [], [], 3.1415, variant)
    assert r0.size == 0 and f0.size == 0


def test_reference_strrt_solutions_valid_on_gpu(pkg, packs, sp):
    """The ST-RRT* solutions the reference logged are valid motions through the C ABI as well."""
    import gzip, json, os
    from conftest import GOLDEN
    with gzip.open(os.path.join(GOLDEN, "reference_logged_paths.json.gz"), "rb") as f:
        logs = json.load(f)
    for key in ("strrt_mass_test", "strrt_flying_boxes", "strrt_multiple_robots"):
        ent = logs[key]
        pk = packs(ent["scene"])
        sc = sp.read_pack(pk)
        dev = pkg.Scene(pk, 0, pkg.PARITY_FIXED if ent["scene"] == "multiple_robots" else pkg.PARITY_REFERENCE_EXACT)
        t, q = np.asarray(ent["t"], np.float64), np.asarray(ent["q"], np.float64)
        keep = t[1:] <= (sc.n_frames - 1) / sc.fps
        for variant in (0, 1):
            res, frac = dev.motion_check_batch(q[:-1][keep], q[1:][keep], t[:-1][keep], t[1:][keep], sc.joint_max_velocity[0], variant)
            assert keep.sum() >= 2 and (res == -1).all() and ((frac > 0) & (frac < 1)).all(), (key, variant, res, frac)


def test_validator_classes_program(pkg, orc, sp, packs, tmp_path):
    """host/OMPLValidators.hpp (MDP::StateValidityCheckerFunctor, MDP::SpaceTimeMotionValidator) driven by
    host/sirrt_validators: single checkMotion calls, the batched form, isValid, clearance and the counters."""
    import os, subprocess
    exe = os.path.join(os.path.dirname(pkg.__file__), "host", "sirrt_validators")
    assert os.path.exists(exe), "host/sirrt_validators not built (python __graft_entry__.py)"
    pk = packs("synB_32_200")
    sc = sp.read_pack(pk)
    ref = orc.Oracle(pk)
    n = 60
    q0, q1, t0, t1 = _random_motions(sp, pk, n, 97, (sc.fps, sc.n_frames))
    (tmp_path / "s.pack").write_bytes(pk)
    with open(tmp_path / "m.txt", "w") as f:
        for e in range(n):
            f.write(" ".join(repr(float(x)) for x in [*q0[e], t0[e], *q1[e], t1[e]]) + "\n")
    out = subprocess.run([exe, str(tmp_path / "s.pack"), str(tmp_path / "m.txt")], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr
    lines = out.stdout.strip().splitlines()
    assert len(lines) == n + 2
    vmax = sc.joint_max_velocity[0]
    codes = [int(x) for x in lines[n].split()[1:]]
    tested = n

    def n_tested(r, e, k):
        if r >= 0:
            return r + 2
        if r != -1:
            return 1
        dq = float(np.sqrt(((q1[e] - q0[e]) ** 2).sum()))
        return max(int(dq / 0.1), int(k * (t1[e] - t0[e]) * sc.fps + 0.5)) + 3

    for e in range(n):
        v0, v1, frac, s2ok, clr = lines[e].split()
        r0, _ = ref.motion_check(q0[e], q1[e], t0[e], t1[e], vmax, 0)
        r1, f1 = ref.motion_check(q0[e], q1[e], t0[e], t1[e], vmax, 1)
        tested += 2 * n_tested(r0, e, 2.0) + n_tested(r1, e, 1.0)
        assert int(v0) == (r0 == -1) and int(v1) == (r1 == -1) and float(frac) == f1 and codes[e] == r0
        assert int(s2ok) == ref.state_valid(q1[e], t1[e])
        fr = int(np.float32(np.float32(t0[e]) * np.float32(sc.fps)))
        col, d = ref.get_distances(q0[e], fr)
        assert abs(float(clr) - (0.0 if col else float(np.float32(d.min())))) < 1e-6
    assert lines[n + 1].split() == ["counters", str(tested), str(3 * n)]


@pytest.mark.parametrize("name", ["flying_boxes", "mass_test", "multiple_robots"])
def test_hull_support_cells_equal_full_scan(pkg, orc, sp, packs, name):
    """GJK support queries through the cube-map candidate cells (csrc/sirrt_tables.hpp: build_hull_cells) pick the
    vertex the full scan picks: rows, cells and per-link distances are identical with the cells switched off, and
    both equal the oracle."""
    pk = packs(name)
    a, b = pkg.Scene(pk), pkg.Scene(pk)
    b.set_option(pkg.OPT_HULL_CELLS, 0)
    q = random_q(sp, pk, 96, 53)
    ta, ba, _ = a.safe_intervals_batch(q, 8)
    tb, bb, _ = b.safe_intervals_batch(q, 8)
    assert np.array_equal(ba, bb) and np.array_equal(ta, tb) and ba.any()
    rt, rb = orc.Oracle(pk).safe_intervals_batch(q, 8)
    assert np.array_equal(ba, rb) and np.array_equal(ta, rt)
    q64 = q.T.astype(np.float64)
    for i in range(0, 96, 8):
        f = (61 * i) % a.n_frames
        ca, da = a.get_distances(q64[i], f)
        cb, db = b.get_distances(q64[i], f)
        assert ca == cb and np.array_equal(da, db)
    b.set_option(pkg.OPT_HULL_CELLS, 1)
    tc, bc, _ = b.safe_intervals_batch(q, 8)
    assert np.array_equal(bc, ba)


def test_concurrent_scenes_are_deterministic(pkg, packs, tmp_path):
    """host/sirrt_stress: 8 host threads, each with its own scene and kNN stores on the same GPU, issue the planner's
    call mix (safe intervals, cell batches, edge tables, kNN add / nearest / radius, stores created and destroyed) on
    identical inputs, in both wait modes; every answer must equal the single-thread answer."""
    import json, os, subprocess
    exe = os.path.join(os.path.dirname(pkg.__file__), "host", "sirrt_stress")
    assert os.path.exists(exe), "host/sirrt_stress not built (python __graft_entry__.py)"
    (tmp_path / "s.pack").write_bytes(packs("mass_test"))
    for mode in ("spin", "yield"):
        r = subprocess.run([exe, str(tmp_path / "s.pack"), "8", "6"], capture_output=True, text=True, timeout=300, env=dict(os.environ, SIRRT_SYNC=mode))
        assert r.returncode == 0, (mode, r.stdout, r.stderr)
        out = json.loads(r.stdout.strip().splitlines()[-1])
        assert out["failed_threads"] == 0 and not any(out["mismatch"].values()), out


@pytest.mark.parametrize("name", ["flying_boxes", "mass_test", "synB_32_200", "multiple_robots"])
def test_windowed_edge_table_equals_full_rows(pkg, sp, packs, name):
    """sirrt_collision_bits_window_f64: inside each row's window (n_tiles 32-frame tiles from first_tile[i]) the words equal
    the full edge-table rows of sirrt_collision_bits_f64 (words outside carry no information).  Windows at the start, in the middle,
    hanging over the end of the scene and entirely past it."""
    pk = packs(name)
    dev = pkg.Scene(pk)
    W = dev.words
    q = random_q(sp, pk, 40, 67).T.astype(np.float64)
    full = dev.collision_bits_f64(q, pkg.SEM_FRAME)
    rng = np.random.default_rng(68)
    for n_tiles in (1, 3):
        ft = rng.integers(-2, W + 2, 40).astype(np.int32)
        ft[:4] = [0, W - 1, W, -5]
        win = dev.collision_bits_window_f64(q, ft, n_tiles, pkg.SEM_FRAME)
        for i in range(40):
            lo = min(max(int(ft[i]), 0), W)
            hi = min(W, lo + n_tiles)
            assert np.array_equal(win[i, lo:hi], full[i, lo:hi]), (name, n_tiles, i, lo, hi)
    assert full.any()


@pytest.mark.parametrize("name,mode", [("flying_boxes", 0), ("mass_test", 0), ("synB_32_200", 0), ("synA_64_300", 0), ("multiple_robots", 0),
                                       ("multiple_robots", 1), ("multiple_robots_cyl", 1)])
def test_slot_split_equals_whole_rows(pkg, orc, sp, packs, name, mode):
    """Planner-sized batches split the dynamic obstacles of a row (k_collide_bits) and the (link, slot) pairs of a cell
    (k_check_cells) over several warps that OR their words / hit flags into zeroed outputs (SIRRT_OPT_SLOT_SPLIT).  Every latency
    entry point must answer what the unsplit launch and the oracle answer -- including repeated single calls, whose internal
    row buffer has to read zero again after each call."""
    pk = packs(name)
    a, b = pkg.Scene(pk, 0, mode), pkg.Scene(pk, 0, mode)
    b.set_option(pkg.OPT_SLOT_SPLIT, 0)
    ref = orc.Oracle(pk, mode)
    q = random_q(sp, pk, 24, 91)
    qd = q.T.astype(np.float64).copy()
    for rep in range(2):  # the second round runs on the buffers the first one left behind
        for qq in qd[:8]:
            ia, ib = a.safe_intervals(qq), b.safe_intervals(qq)
            assert ia == ib == ref.safe_intervals(qq), (name, rep)
    for sem in (pkg.SEM_FRAME, pkg.SEM_INTERVAL):
        fa, fb = a.collision_bits_f64(qd, sem), b.collision_bits_f64(qd, sem)
        assert np.array_equal(fa, fb), (name, sem)
    ta, ba, _ = a.safe_intervals_batch(q, 8)
    tb, bb, _ = b.safe_intervals_batch(q, 8)
    assert np.array_equal(ba, bb) and np.array_equal(ta, tb)
    assert np.array_equal(ba, a.collision_bits_f64(qd, pkg.SEM_INTERVAL))
    W = a.words
    ft = ((np.arange(24) * 5) % max(W, 1)).astype(np.int32)
    wa, wb = a.collision_bits_window_f64(qd, ft, 3, pkg.SEM_FRAME), b.collision_bits_window_f64(qd, ft, 3, pkg.SEM_FRAME)
    full = b.collision_bits_f64(qd, pkg.SEM_FRAME)
    for i in range(24):
        lo, hi = int(ft[i]), min(W, int(ft[i]) + 3)
        assert np.array_equal(wa[i, lo:hi], wb[i, lo:hi]) and np.array_equal(wa[i, lo:hi], full[i, lo:hi]), (name, i)
    fr = ((np.arange(24, dtype=np.int64) * 131) % ref.n_frames).astype(np.int32)
    fr[3], fr[7] = -1, ref.n_frames  # out-of-range frames answer "collided"
    ca, cb = a.check_frames_batch_f64(qd, fr), b.check_frames_batch_f64(qd, fr)
    assert np.array_equal(ca, cb)
    want = [True if (f < 0 or f >= ref.n_frames) else bool(ref.check_frame(qq, int(f))) for qq, f in zip(qd, fr)]
    assert [bool(x) for x in ca] == want
    for rep in range(2):
        for qq, f in zip(qd[:6], fr[:6]):
            if 0 <= f < ref.n_frames:
                assert a.check_frame(qq, int(f)) == b.check_frame(qq, int(f)) == bool(ref.check_frame(qq, int(f)))
    # the same calls with the kernels of a call chained as programmatic dependent launches
    c = pkg.Scene(pk, 0, mode)
    c.set_option(pkg.OPT_PDL, 1)
    c.set_option(pkg.OPT_INLINE_RECHECK, 0)  # ... and with the FP64 re-check kernel always queued
    for qq, f in zip(qd[:8], fr[:8]):
        assert c.safe_intervals(qq) == a.safe_intervals(qq)
        if 0 <= f < ref.n_frames:
            assert c.check_frame(qq, int(f)) == a.check_frame(qq, int(f))
    assert np.array_equal(c.check_frames_batch_f64(qd, fr), ca)
    # every FP64 item handed back to the host (the CONTINUE path of the completion flag)
    d = pkg.Scene(pk, 0, mode)
    d.set_option(pkg.OPT_INLINE_RECHECK, 1)
    for qq, f in zip(qd, fr):
        assert d.safe_intervals(qq) == a.safe_intervals(qq)
    assert np.array_equal(d.check_frames_batch_f64(qd, fr), ca)
    # the FP64 re-check started from scratch instead of from the FP32 run's simplex
    e = pkg.Scene(pk, 0, mode)
    e.set_option(pkg.OPT_SEED_RECHECK, 0)
    for qq in qd:
        assert e.safe_intervals(qq) == a.safe_intervals(qq)
    assert np.array_equal(e.check_frames_batch_f64(qd, fr), ca)
    assert np.array_equal(e.collision_bits_f64(qd, pkg.SEM_INTERVAL), a.collision_bits_f64(qd, pkg.SEM_INTERVAL))
    # hull pair tests without the oriented-box cull in front of them
    g = pkg.Scene(pk, 0, mode)
    g.set_option(pkg.OPT_OBB_CULL, 0)
    for qq in qd[:12]:
        assert g.safe_intervals(qq) == a.safe_intervals(qq)
    assert np.array_equal(g.check_frames_batch_f64(qd, fr), ca)
    assert np.array_equal(g.collision_bits_f64(qd, pkg.SEM_FRAME), a.collision_bits_f64(qd, pkg.SEM_FRAME))
    tg, bg, _ = g.safe_intervals_batch(q, 8)
    assert np.array_equal(bg, ba) and np.array_equal(tg, ta)
    wc = c.collision_bits_window_f64(qd, ft, 3, pkg.SEM_FRAME)
    for i in range(24):
        lo, hi = int(ft[i]), min(W, int(ft[i]) + 3)
        assert np.array_equal(wc[i, lo:hi], wa[i, lo:hi]), (name, i)


@pytest.mark.parametrize("name", ["flying_boxes", "synA_32_200", "multiple_robots"])
def test_warp_fk_agrees_with_the_sequential_chain(pkg, orc, sp, packs, name):
    """The planner-sized kernels build their link table from a warp-parallel FK (parallel prefix over the joints,
    fk_link_of_lane); the products associate differently from the sequential chain of fk_links (the one the FP64 re-check and
    the oracle walk, RobotObstacleFCL.cpp:287-386), so the transforms agree to rounding, not bit for bit: 1e-13 here, against
    the FP32 link table's 6e-8 and the enclosure certificate's 1e-9 m margin."""
    pk = packs(name)
    dev = pkg.Scene(pk)
    ref = orc.Oracle(pk)
    q = random_q(sp, pk, 64, 5).T.astype(np.float64)
    worst = 0.0
    for qq in q:
        R1, t1 = dev.fk(qq)
        R2, t2 = dev.fk_warp(qq)
        worst = max(worst, float(np.abs(R1 - R2).max()), float(np.abs(t1 - t2).max()))
        Ro, to = ref.fk(qq)
        assert np.abs(R2 - Ro).max() < 1e-12 and np.abs(t2 - to).max() < 1e-12
    assert worst < 1e-13, worst


@pytest.mark.gpu
def test_knn_radius_begin_end_equals_radius(pkg):
    """sirrt_knn_radius_begin / _end (the planner launches the radius query of set_parent before its safe-interval call and
    collects it afterwards) answer what sirrt_knn_radius answers; a pending query is settled and dropped by any other call
    on the store; _end without _begin is an error; more hits than the pinned block carries (1023) still arrive."""
    rng = np.random.default_rng(77)
    dim = 6
    store = pkg.KnnIndex(dim, 64)
    idx0, n0 = store.radius_end()          # empty store: 0 hits, no query needed
    assert n0 == 0
    pts = rng.uniform(-1.0, 1.0, (3000, dim))
    store.add(pts)
    with pytest.raises(pkg.SirrtError):
        store.radius_end()
    for r2 in (0.05, 0.6, 30.0):           # 30.0: every point, beyond the 1023 hits of the pinned block
        q = rng.uniform(-1.0, 1.0, dim)
        want, n_want = store.radius(q, r2)
        store.radius_begin(q, r2)
        got, n_got = store.radius_end()
        assert n_got == n_want and np.array_equal(got, want)
    q = rng.uniform(-1.0, 1.0, dim)
    store.radius_begin(q, 0.6)             # left pending: the next calls settle it
    i1, d1 = store.nearest(q[None, :])
    assert i1[0] == int(np.argmin(((pts - q) ** 2).sum(1)))
    store.radius_begin(q, 0.6)
    store.add(rng.uniform(-1.0, 1.0, (1, dim)))
    want, n_want = store.radius(q, 0.6)
    store.radius_begin(q, 0.6)
    store.radius_begin(q, 0.7)             # a second begin drops the first
    got, n_got = store.radius_end()
    w2, n2 = store.radius(q, 0.7)
    assert n_got == n2 and np.array_equal(got, w2)
    store.close()
