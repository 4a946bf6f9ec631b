"""CPU model of the k-NN tile kernel's work (no GPU): candidates per tile, lanes covered / deferred, for a
given cloud, k and cell size.  Guides design choices between GPU runs.
usage: sim_tile.py [n] [k] [cell] """
import sys, numpy as np
from scipy.spatial import cKDTree
sys.path.insert(0, ".")
from pointcloudprocessing_b200 import synth

def grid_sort(p, cell, shift=0.0):
    lo = p[:, :3].min(0).astype(np.float64); hi = p[:, :3].max(0).astype(np.float64)
    ext = hi - lo
    order = np.argsort(-ext, kind="stable")
    ax = [order[1], order[2], order[0]]  # ax0 fastest = 2nd longest, ax1 = shortest, ax2 = longest
    dim = [int(np.floor(ext[a] / cell)) + 1 for a in ax]
    t = [(p[:, ax[a]].astype(np.float64) - lo[ax[a]]) / cell for a in range(3)]
    c = [np.clip(np.floor(t[a]).astype(np.int64), 0, dim[a] - 1) for a in range(3)]
    f = [np.clip(t[a] - c[a], 0, 1) for a in range(3)]
    cid = (c[2] * dim[1] + c[1]) * dim[0] + c[0]
    perm = np.argsort(cid, kind="stable")
    return ax, dim, [x[perm] for x in c], [x[perm] for x in f], cid[perm], perm

def model(p, k, cell, CAP=32, verbose=True):
    n = p.shape[0]
    ax, dim, c, f, cid, perm = grid_sort(p, cell)
    ncell = dim[0] * dim[1] * dim[2]
    E = np.zeros(ncell + 1, np.int64)
    np.add.at(E, cid + 1, 1); E = np.cumsum(E)
    tree = cKDTree(p[:, :3].astype(np.float64))
    dk = tree.query(p[perm, :3].astype(np.float64), k=k, workers=-1)[0][:, k - 1]
    nt = (n + 31) // 32
    pad = nt * 32 - n
    def P(x, fill): return np.concatenate([x, np.full(pad, fill, x.dtype)]).reshape(nt, 32)
    row = cid // dim[0]
    rowT = P(row, -1); c0T = P(c[0], 0); validT = P(np.ones(n, bool), False)
    # majority row per tile = mode; approximate by the row of the median lane, then the most common of first/last
    first, last = rowT[:, 0], rowT[:, 31]
    cnt_first = (rowT == first[:, None]).sum(1); cnt_last = (rowT == last[:, None]).sum(1)
    row0 = np.where(cnt_first >= cnt_last, first, last)
    in_grp = (rowT == row0[:, None]) & validT
    c0min = np.where(in_grp, c0T, 1 << 30).min(1); c0max = np.where(in_grp, c0T, -1).max(1)
    c1 = row0 % dim[1]; c2 = row0 // dim[1]
    R = 1
    xlo = np.maximum(c0min - R, 0); xhi = np.minimum(c0max + R, dim[0] - 1)
    total = np.zeros(nt, np.int64)
    for d2 in (-1, 0, 1):
        for d1 in (-1, 0, 1):
            e1 = c1 + d1; e2 = c2 + d2
            ok = (e1 >= 0) & (e1 < dim[1]) & (e2 >= 0) & (e2 < dim[2])
            base = (np.clip(e2, 0, dim[2] - 1) * dim[1] + np.clip(e1, 0, dim[1] - 1)) * dim[0]
            total += np.where(ok, E[base + xhi + 1] - E[base + xlo], 0)
    f0T, f1T, f2T = P(f[0], 0.5), P(f[1], 0.5), P(f[2], 0.5)
    inf = 1e18
    m0a = np.where((xlo > 0)[:, None], (c0T - xlo[:, None]) + f0T, inf)
    m0b = np.where((xhi < dim[0] - 1)[:, None], (xhi[:, None] + 1 - c0T) - f0T, inf)
    m1a = np.where((c1 - R > 0)[:, None], R + f1T, inf); m1b = np.where((c1 + R < dim[1] - 1)[:, None], R + 1 - f1T, inf)
    m2a = np.where((c2 - R > 0)[:, None], R + f2T, inf); m2b = np.where((c2 + R < dim[2] - 1)[:, None], R + 1 - f2T, inf)
    cover = np.minimum.reduce([m0a, m0b, m1a, m1b, m2a, m2b]) * cell
    dkT = P(dk, 0.0)
    covered = in_grp & (dkT < cover)
    unc = in_grp & ~covered
    offrow = validT & ~in_grp
    res = dict(cell=cell, dim=dim, cand_per_tile=total.mean(), uncovered=unc.sum() / n, offrow=offrow.sum() / n,
               deferred=(unc.sum() + offrow.sum()) / n, unc_per_tile_hist=np.bincount(unc.sum(1), minlength=33)[:33],
               span=(c0max - c0min + 1).mean())
    if verbose:
        print(f"cell {cell:.3f} dim {dim} cand/tile {res['cand_per_tile']:.0f} span {res['span']:.2f} uncovered {100*res['uncovered']:.2f}% "
              f"off-row {100*res['offrow']:.2f}% deferred {100*res['deferred']:.2f}%")
        h = res['unc_per_tile_hist']
        print("  tiles by #uncovered lanes: 0:%d 1:%d 2:%d 3-4:%d 5-8:%d 9-16:%d 17-32:%d" % (h[0], h[1], h[2], h[3:5].sum(), h[5:9].sum(), h[9:17].sum(), h[17:].sum()))
    res.update(dk=dk, cover=cover, in_grp=in_grp, c=c, f=f, cid=cid, perm=perm, E=E, dimv=dim, total=total, unc=unc)
    return res

if __name__ == "__main__":
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
    k = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    p = synth.urban(3, n, extent=200.0 * (n / 1e6) ** 0.5, order="random")
    cells = [float(sys.argv[3])] if len(sys.argv) > 3 else [0.8, 0.9, 1.0, 1.1, 1.264, 1.4]
    for cell in cells:
        model(p, k, cell)
