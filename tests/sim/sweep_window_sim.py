"""Design-time simulation (NumPy, CPU) of the k_sweep v5 B-stage: analytic frame windows per (link, tile, slot)
survivor and the tight per-frame bounds.  Not a test; prints stage counts used in DESIGN.md and checks that the
window is a superset of the frames the per-frame cull keeps.  usage: python tests/sim/sweep_window_sim.py [n_cfg]"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import __graft_entry__ as g

def seg_box_dist(a, b, h, iters=70):
    # a,b [...,3] segment end points in the box frame; h [...,3]; golden-section on convex g(t)
    d = b - a
    lo = np.zeros(a.shape[:-1]); hi = np.ones(a.shape[:-1])
    gr = (np.sqrt(5) - 1) / 2
    def gfun(t):
        p = a + t[..., None] * d
        e = np.maximum(np.abs(p) - h, 0)
        return (e * e).sum(-1)
    x1 = hi - gr * (hi - lo); x2 = lo + gr * (hi - lo)
    f1 = gfun(x1); f2 = gfun(x2)
    for _ in range(iters):
        m = f1 < f2
        hi = np.where(m, x2, hi); lo = np.where(m, lo, x1)
        x1n = hi - gr * (hi - lo); x2n = lo + gr * (hi - lo)
        x1, x2 = x1n, x2n
        f1 = gfun(x1); f2 = gfun(x2)
    return np.sqrt(np.minimum(np.minimum(f1, f2), np.minimum(gfun(np.zeros_like(lo)), gfun(np.ones_like(lo)))))

def main():
    ncfg = int(sys.argv[1]) if len(sys.argv) > 1 else 32
    sp, orc = g.load_tools(), g.load_oracle()
    scene = sp.synthetic_scene(64, 1000, 30, 42, "A")
    pack = sp.write_pack(scene)
    O = orc.Oracle(pack)
    q = sp.random_configs(scene.robot, ncfg, 42)
    T = 1000; C = 64; L = O.n_links
    shapes = scene.robot.shapes[2:]
    rl = np.array([s.radius for s in shapes]); hl = np.array([s.length / 2 for s in shapes])
    ctr = np.stack([o.poses[:, :3].astype(np.float64) for o in scene.obstacles])  # [C,T,3]
    h = np.array([o.dims[0] / 2 for o in scene.obstacles], np.float64)            # [C]
    band = 2e-5
    tot = dict(pairs=0, b1=0, in_=0, hit0=0, shell=0, thit=0, tfree=0, und=0, und_hit=0, win=0, win_miss=0, win_empty=0, cells_hit=0)
    for c in range(ncfg):
        R, t = O.fk(q[:, c].astype(np.float64))
        m = t; u = R[:, :, 2]                          # [L,3]
        a = m - u * hl[:, None]; b = m + u * hl[:, None]
        ab = b - a
        # point-segment distance for all (l, j, f)
        w = ctr[None] - a[:, None, None, :]            # [L,C,T,3]
        den = (ab * ab).sum(-1)
        tt = np.clip((w * ab[:, None, None, :]).sum(-1) / den[:, None, None], 0, 1)
        e = w - tt[..., None] * ab[:, None, None, :]   # from closest axis point to the centre
        dps = np.sqrt((e * e).sum(-1))
        hi = rl[:, None, None] + h[None, :, None] * np.sqrt(3) + band
        lo = rl[:, None, None] + h[None, :, None] - band
        inn = dps <= hi; hit0 = inn & (dps < lo); shell = inn & ~hit0
        # tight bounds (identity rotation: world == box frame)
        ae = np.abs(e)
        qq = np.maximum(ae - h[None, :, None, None], 0)
        dbox = np.sqrt((qq * qq).sum(-1))
        thit = shell & (dbox < rl[:, None, None] - band)
        sup = (ae * h[None, :, None, None]).sum(-1) / np.maximum(dps, 1e-30)
        tfree = shell & ~thit & (dps - sup > rl[:, None, None] + band)
        und = shell & ~thit & ~tfree
        idx = np.nonzero(und)
        la = a[idx[0]] - ctr[idx[1], idx[2]]; lb = b[idx[0]] - ctr[idx[1], idx[2]]
        dex = seg_box_dist(la, lb, np.repeat(h[idx[1]][:, None], 3, 1)) - rl[idx[0]]
        # sanity of the tight bounds against the exact distance on a sample of the shell
        ish = np.nonzero(thit | tfree)
        if len(ish[0]):
            sel = np.random.default_rng(c).choice(len(ish[0]), min(2000, len(ish[0])), replace=False)
            i0, i1, i2 = ish[0][sel], ish[1][sel], ish[2][sel]
            dd = seg_box_dist(a[i0] - ctr[i1, i2], b[i0] - ctr[i1, i2], np.repeat(h[i1][:, None], 3, 1)) - rl[i0]
            th = thit[i0, i1, i2]
            assert (dd[th] < 1e-9).all(), "tight HIT bound wrong"
            assert (dd[~th] > -1e-9).all(), "tight FREE bound wrong"
        tot["pairs"] += L * C * 32
        for k, v in (("in_", inn), ("hit0", hit0), ("shell", shell), ("thit", thit), ("tfree", tfree), ("und", und)):
            tot[k] += int(v.sum())
        tot["und_hit"] += int((dex < 0).sum())
        # analytic windows per (l, tile, j): two-piece linear model from the true centres
        for tile in range(32):
            f0 = tile * 32; nf = min(32, T - f0)
            cc = ctr[:, f0:f0 + nf]                                        # [C,nf,3]
            fr = np.arange(nf)
            # B1: swept capsule (segment between first/last centre + deviation)
            # model: best split
            best_dev = np.full(C, np.inf); Pa = np.zeros((C, 3)); Va = np.zeros((C, 3)); Pb = np.zeros((C, 3)); Vb = np.zeros((C, 3)); fs = np.full(C, nf)
            for s in range(1, nf + 1):
                pa = cc[:, 0]; va = (cc[:, s - 1] - cc[:, 0]) / max(s - 1, 1)
                moda = pa[:, None] + va[:, None] * fr[None, :, None]
                if s < nf:
                    pb = cc[:, s]; vb = (cc[:, nf - 1] - cc[:, s]) / max(nf - 1 - s, 1)
                    modb = pb[:, None] + vb[:, None] * (fr[None, :, None] - s)
                else:
                    pb = pa; vb = va; modb = moda
                mod = np.where((fr < s)[None, :, None], moda, modb)
                dev = np.sqrt(((mod - cc) ** 2).sum(-1)).max(1)
                better = dev < best_dev
                best_dev = np.where(better, dev, best_dev)
                for arr, val in ((Pa, pa), (Va, va), (Pb, pb), (Vb, vb)):
                    arr[better] = val[better]
                fs[better] = s
            inn_t = inn[:, :, f0:f0 + nf]                                   # [L,C,nf]
            surv = inn_t.any(-1)                                            # ideal B1 survivors (lower bound on real B1)
            for l in range(L):
                Rl = rl[l] + h * np.sqrt(3) + band + best_dev + 2e-4        # [C]
                Rc = Rl + hl[l]
                inv = 1.0 / den[l]
                mask = np.zeros((C, nf), bool)
                for (P, V, g0, g1) in ((Pa, Va, np.zeros(C), fs - 1), (Pb, Vb, fs.astype(float), np.full(C, nf - 1.0))):
                    w0 = P - m[l] - V * 0  # at local g=0 -> frame g0 ... for piece b P is at frame fs
                    A1 = (V * V).sum(-1); B1 = (w0 * V).sum(-1); C1 = (w0 * w0).sum(-1)
                    wa = (w0 * ab[l]).sum(-1); va_ = (V * ab[l]).sum(-1)
                    A2 = A1 - va_ * va_ * inv; B2 = B1 - wa * va_ * inv; C2 = C1 - wa * wa * inv
                    def win(A, B, Cc, R2):
                        disc = B * B - A * (Cc - R2)
                        ok = disc >= 0
                        sq = np.sqrt(np.maximum(disc, 0))
                        small = A < 1e-12
                        glo = np.where(small, -1e9, (-B - sq) / np.where(small, 1, A))
                        ghi = np.where(small, 1e9, (-B + sq) / np.where(small, 1, A))
                        ok = np.where(small, Cc <= R2, ok)
                        return ok, glo, ghi
                    ok1, l1, h1 = win(A1, B1, C1, Rc * Rc)
                    ok2, l2, h2 = win(A2, B2, C2, Rl * Rl)
                    glo = np.floor(np.maximum(l1, l2)) - 1; ghi = np.ceil(np.minimum(h1, h2)) + 1
                    gl = fr[None, :] - g0[:, None]
                    pm = (ok1 & ok2)[:, None] & (gl >= glo[:, None]) & (gl <= ghi[:, None]) & (fr[None, :] >= g0[:, None]) & (fr[None, :] <= g1[:, None])
                    mask |= pm
                tot["win"] += int(mask[surv[l]].sum())
                tot["win_empty"] += int((~mask.any(-1) & surv[l]).sum())
                miss = inn_t[l] & ~mask
                tot["win_miss"] += int(miss.sum())
                tot["b1"] += int(surv[l].sum())
    n = ncfg
    print({k: v / n for k, v in tot.items()})

main()
