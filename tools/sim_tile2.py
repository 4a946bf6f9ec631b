"""Model: per-tile stencil half-width chosen from the tile's span (cells of the majority row the 32 queries touch)."""
import sys, numpy as np
sys.path.insert(0, "."); sys.path.insert(0, "tools")
from pointcloudprocessing_b200 import synth
import sim_tile as S

def eval_rule(res, cell, rule, pair_cost=18.0, wl_cost=1750.0, fixed=5500.0, maxcand=None):
    dim = res['dimv']; E = res['E']; c = res['c']; f = res['f']; cid = res['cid']; dk = res['dk']
    n = cid.shape[0]; nt = (n + 31) // 32; pad = nt * 32 - n
    def P(x, fill): return np.concatenate([x, np.full(pad, fill, x.dtype)]).reshape(nt, 32)
    row = cid // dim[0]
    rowT = P(row, -1); c0T = P(c[0], 0); validT = P(np.ones(n, bool), False)
    first, last = rowT[:, 0], rowT[:, 31]
    cnt_first = (rowT == first[:, None]).sum(1); cnt_last = (rowT == last[:, None]).sum(1)
    row0 = np.where(cnt_first >= cnt_last, first, last)
    in_grp = (rowT == row0[:, None]) & validT
    c0min = np.where(in_grp, c0T, 1 << 30).min(1); c0max = np.where(in_grp, c0T, -1).max(1)
    span = c0max - c0min + 1
    c1 = row0 % dim[1]; c2 = row0 // dim[1]
    Rt = rule(span, in_grp.sum(1))
    xlo = np.maximum(c0min - Rt, 0); xhi = np.minimum(c0max + Rt, dim[0] - 1)
    total = np.zeros(nt, np.int64)
    for d2 in range(-3, 4):
        for d1 in range(-3, 4):
            use = (np.abs(d2) <= Rt) & (np.abs(d1) <= Rt)
            e1 = c1 + d1; e2 = c2 + d2
            ok = use & (e1 >= 0) & (e1 < dim[1]) & (e2 >= 0) & (e2 < dim[2])
            base = (np.clip(e2, 0, dim[2] - 1) * dim[1] + np.clip(e1, 0, dim[1] - 1)) * dim[0]
            total += np.where(ok, E[base + xhi + 1] - E[base + xlo], 0)
    f0T, f1T, f2T = P(f[0], 0.5), P(f[1], 0.5), P(f[2], 0.5)
    inf = 1e18; R = Rt[:, None]
    m0a = np.where((xlo > 0)[:, None], (c0T - xlo[:, None]) + f0T, inf)
    m0b = np.where((xhi < dim[0] - 1)[:, None], (xhi[:, None] + 1 - c0T) - f0T, inf)
    m1a = np.where((c1[:, None] - R > 0), R + f1T, inf); m1b = np.where((c1[:, None] + R < dim[1] - 1), R + 1 - f1T, inf)
    m2a = np.where((c2[:, None] - R > 0), R + f2T, inf); m2b = np.where((c2[:, None] + R < dim[2] - 1), R + 1 - f2T, inf)
    cover = np.minimum.reduce([m0a, m0b, m1a, m1b, m2a, m2b]) * cell
    dkT = P(dk, 0.0)
    covered = in_grp & (dkT < cover)
    unc = in_grp & ~covered
    offrow = validT & ~in_grp
    if maxcand is not None:  # tiles above the cap fall back to R = 1 ... (not modelled)
        pass
    deferred = unc.sum() + offrow.sum()
    cost = (total * pair_cost).sum() + deferred * wl_cost + nt * fixed
    hist = [int((Rt == r).sum()) for r in (1, 2, 3)]
    print(f"  tiles R1/R2/R3 {hist}  cand/tile {total.mean():.0f} (R1 {total[Rt==1].mean():.0f}, R2 {total[Rt==2].mean() if hist[1] else 0:.0f}, R3 {total[Rt==3].mean() if hist[2] else 0:.0f})"
          f"  uncovered {100*unc.sum()/n:.2f}% deferred {100*deferred/n:.2f}%  model cost/tile {cost/nt/1000:.1f}k")
    return cost / nt

if __name__ == "__main__":
    n = 1_000_000; k = int(sys.argv[1]) if len(sys.argv) > 1 else 20
    p = synth.urban(3, n, extent=200.0, order="random")
    for cell in [0.9, 1.0, 1.1, 1.264]:
        res = S.model(p, k, cell, verbose=True)
        for (s1, s2) in [(99, 99), (6, 99), (4, 99), (8, 99), (6, 16), (4, 12), (3, 8)]:
            print(f" rule span<={s1}:R1, <={s2}:R2 else R3")
            eval_rule(res, cell, lambda span, cnt: np.where(span <= s1, 1, np.where(span <= s2, 2, 3)))
