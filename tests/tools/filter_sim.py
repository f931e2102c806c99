"""CPU analysis tool (not a test, not product code): pass rates of the tile kernel's block filter and of tighter variants on a synthetic series,
with the final profile as the thresholds.  usage: python tests/tools/filter_sim.py [n] [m] [walk|uniform]"""
import sys, os, numpy as np, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', '..', 'oracle'))   # the oracle is test infrastructure: this analysis tool lives under tests/
from oracle_py import Oracle, random_walk
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 16
m = int(sys.argv[2]) if len(sys.argv) > 2 else 256
kind = sys.argv[3] if len(sys.argv) > 3 else 'walk'
rng = np.random.default_rng(1234)
T = rng.random(n) if kind == 'uniform' else np.cumsum(rng.standard_normal(n))
o = Oracle(fast=True)
t0 = time.time()
MP, MPI, corr, _ = o.run(T, m, 4096, nthreads=8)
print('oracle', time.time() - t0, 's; median best', np.median(corr), 'p10', np.percentile(corr, 10))
inp = o.precompute(T, m)
df, dg, inv, qt0 = inp['df'], inp['dg'], inp['inv'], inp['QT']
sublen = n - m + 1
excl = m // 4
norm = 1.0 / inv
best = corr
PADN = 600
def pad(a, v): return np.concatenate([a, np.full(PADN, v)])
dfp, dgp, normp, bestp = pad(df, 0), pad(dg, 0), pad(norm, np.inf), pad(best, 1.0)
res = {}
def acc(k, lane_pass, warp_pass, nl, nw):
    r = res.setdefault(k, [0, 0, 0, 0]); r[0] += lane_pass; r[1] += nl; r[2] += warp_pass; r[3] += nw
bands = [int(x) for x in np.linspace(0, (sublen - excl) // 256 - 2, 24)]
for wb in bands:
    d0 = excl + wb * 256
    R = (sublen - d0) // 8 * 8
    if R < 64: continue
    r = np.arange(R)
    # QT[r, j] along diagonal d0+j: QT[0] = qt0[d0+j]; QT[r] = QT[r-1] + df[r] dg[c] + df[c] dg[r]
    c = r[:, None] + d0 + np.arange(256)[None, :]
    cc = np.minimum(c, sublen + PADN - 1)
    upd = df[r][:, None] * dgp[cc] + dfp[cc] * dg[r][:, None]
    upd[0] = 0
    Q = np.cumsum(upd, axis=0) + pad(qt0, 0)[d0:d0 + 256][None, :]
    valid = c < sublen
    Q = np.where(valid, Q, -np.inf)
    Nr, Br = norm[r], best[r]
    NC, BC = normp[cc], bestp[cc]
    thr_exact = Nr[:, None] * NC * np.minimum(Br[:, None], BC)
    pe = Q >= thr_exact
    B = R // 8
    pe_blk = pe.reshape(B, 8, 32, 8).any(axis=(1, 3))
    acc('exact(any cell improves)', pe_blk.sum(), pe_blk.any(axis=1).sum(), pe_blk.size, B)
    Qb = Q.reshape(B, 8, 32, 8)
    qmax = Qb.max(axis=(1, 3))                       # [B, 32]
    # current: rows [8b, 8b+8), columns window [y, y+15), y = 8b + d0 + 8l
    An = (Br * Nr).reshape(B, 8).min(axis=1); Nn = Nr.reshape(B, 8).min(axis=1)
    y = (8 * np.arange(B))[:, None] + d0 + 8 * np.arange(32)[None, :]
    win = y[:, :, None] + np.arange(15)[None, None, :]
    winc = np.minimum(win, sublen + PADN - 1)
    ncw = normp[winc]; bcw = (bestp * normp)[winc]
    theta = np.minimum(An[:, None] * ncw.min(axis=2), Nn[:, None] * bcw.min(axis=2))
    p = qmax >= theta * (1 - 2.0**-18)
    acc('current 8x8 block', p.sum(), p.any(axis=1).sum(), p.size, B)
    # RB = 4 half blocks: rows [4h,4h+4) columns [y+4h, y+4h+11)
    for hname, nh in (('half blocks 4x8', 2), ('quarter blocks 2x8', 4), ('per row 1x8', 8)):
        rh = 8 // nh
        ph = np.zeros((B, 32), bool)
        for h in range(nh):
            qm = Qb[:, h * rh:(h + 1) * rh].max(axis=(1, 3))
            A_h = (Br * Nr).reshape(B, 8)[:, h * rh:(h + 1) * rh].min(axis=1); N_h = Nr.reshape(B, 8)[:, h * rh:(h + 1) * rh].min(axis=1)
            w0, w1 = h * rh, h * rh + rh + 7
            th = np.minimum(A_h[:, None] * ncw[:, :, w0:w1].min(axis=2), N_h[:, None] * bcw[:, :, w0:w1].min(axis=2))
            ph |= qm >= th * (1 - 2.0**-18)
        acc(hname, ph.sum(), ph.any(axis=1).sum(), ph.size, B)

    # stage-2 candidates inside a replayed block (condition: block passed the current filter in some lane)
    rep = p.any(axis=1)                                   # [B] warp-blocks replayed
    Qr = Qb.max(axis=3)                                   # [B, 8, 32] row max per lane
    NCb = NC.reshape(B, 8, 32, 8); BCb = BC.reshape(B, 8, 32, 8)
    Nr8 = Nr.reshape(B, 8); Br8 = Br.reshape(B, 8)
    # (a) per row 1x8, product-of-minima form
    th_a = np.minimum((Br8 * Nr8)[:, :, None] * NCb.min(axis=3), Nr8[:, :, None] * (BCb * NCb).min(axis=3))
    pa = Qr >= th_a * (1 - 2.0**-18)
    # (b) row P-bound: QTmax * inv_r * max inv_c >= min(best_r, min best_c)
    th_b = Nr8[:, :, None] * NCb.min(axis=3) * np.minimum(Br8[:, :, None], BCb.min(axis=3))
    pb = Qr >= th_b * (1 - 2.0**-18)
    pex = pe.reshape(B, 8, 32, 8).any(axis=3)
    for nm_, pp in (('row 1x8 prod-of-min', pa), ('row P-bound', pb), ('row exact', pex)):
        r2 = res.setdefault('stage2 ' + nm_, [0, 0, 0, 0])
        r2[0] += pp[rep].sum(); r2[1] += pp[rep].size; r2[2] += pp[rep].any(axis=2).sum(); r2[3] += pp[rep].any(axis=2).size
    # slack variants on the current filter: how much tighter would it have to be
    for s in (0.99, 0.98, 0.95, 0.9):
        p2 = qmax * s >= theta
        acc(f'current with QT*{s}', p2.sum(), p2.any(axis=1).sum(), p2.size, B)
for k, (lp, nl, wp, nw) in res.items():
    print(f'{k:32s} lane-blocks passing {100 * lp / nl:7.3f} %   warp-blocks replayed {100 * wp / nw:7.3f} %')
