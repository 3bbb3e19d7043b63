import os, sys, numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import __graft_entry__ as g
import oracle_ref
pkg = g.load_package(); L = pkg._lib; tg = pkg.tablegen
ref = oracle_ref.load()
raw = tg.vinet_sp5_tables()
r = ref.SpinerRhoT(raw)
n = 1 << 16
rho, T = pkg.workloads.spiner_inputs(n, seeds=(5, 6))
sie = r.eval(L.E_FROM_RHO_T, rho, T, nthreads=0)
lam = np.full((n, 2), np.nan)
T_ref = r.eval(L.T_FROM_RHO_E, rho, sie, lam, nthreads=0)
lT_ref = lam[:, 1].copy()
tab = r.get_table("sie").reshape(256, 256)
lTMin, lTMax = r.get_scalar("lTMin"), r.get_scalar("lTMax")
gR, gT = raw.struct.lrho, raw.struct.lT
EPS = 10 * np.finfo(float).eps
def to_log(x, off): return tg.fast_log10(np.abs(np.maximum(x, -off) + off) + EPS)
def weights(gr, x):
    idx = 1.0 / gr.dx[0]
    i = np.clip(np.trunc(idx * (x - gr.min[0])).astype(np.int64), 0, gr.npoints[0] - 2)
    fx = i * gr.dx[0] + gr.min[0]
    w1 = idx * (x - fx); return i, 1.0 - w1, w1
lRho = to_log(rho, raw.struct.rho_offset)
ix2, w20, w21 = weights(gR, lRho)
def f(x, m):
    i, w10, w11 = weights(gT, x)
    j = ix2[m]
    return (w20[m] * (w10 * tab[j, i] + w11 * tab[j, i + 1]) + w21[m] * (w10 * tab[j + 1, i] + w11 * tab[j + 1, i + 1])) - sie[m]
allm = np.arange(n)
a = np.full(n, lTMin); b = np.full(n, lTMax); guess = np.full(n, 0.5 * (lTMin + lTMax))
ya = f(a, allm); yg = f(guess, allm); yb = np.zeros(n)
c1 = ya * yg <= 0
b = np.where(c1, guess, b); yb = np.where(c1, yg, yb)
ybb = f(np.full(n, lTMax), allm)
c2 = ~c1 & (yg * ybb <= 0)
yb = np.where(~c1, ybb, yb)
a = np.where(c2, guess, a); ya = np.where(c2, yg, ya)
assert (c1 | c2).all()
sign = np.where(ya < 0, 1.0, -1.0); ya *= sign; yb *= sign
b1 = np.zeros(n, int); b2 = np.zeros(n, int); it = np.zeros(n, int); nev = np.zeros(n, int)
seg_evals_same = np.zeros(n, int)
active = np.ones(n, bool)
with np.errstate(all="ignore"):
  for step in range(2000):
    active = (b - a > 2e-14) & ((np.abs(ya) > 1e-14) | (np.abs(yb) > 1e-14)) & (it < 1000)
    if not active.any(): break
    m = np.nonzero(active)[0]
    c = (a[m] * yb[m] - b[m] * ya[m]) / (yb[m] - ya[m])
    ea = c == a[m]; eb = (c == b[m]) & ~ea
    b[m[ea]] = a[m[ea]]; a[m[eb]] = b[m[eb]]
    k = ~(ea | eb); mk = m[k]; ck = c[k]
    yc = sign[mk] * f(ck, mk); nev[mk] += 1
    pos = yc > 0; neg = yc < 0; zer = ~(pos | neg)
    p = mk[pos]; b[p] = ck[pos]; yb[p] = yc[pos]; b1[p] += 1; ya[p] *= np.where(b1[p] > 1, 0.5, 1.0); b2[p] = 0
    q = mk[neg]; a[q] = ck[neg]; ya[q] = yc[neg]; b2[q] += 1; yb[q] *= np.where(b2[q] > 1, 0.5, 1.0); b1[q] = 0
    z = mk[zer]; a[z] = ck[zer]; b[z] = ck[zer]
    it[mk] += 1
lT = 0.5 * (a + b)
print("bit-identical lT vs oracle: %.5f" % (lT.view(np.uint64) == lT_ref.view(np.uint64)).mean())
print("iterations: mean %.2f median %d p90 %d p99 %d max %d" % (it.mean(), np.median(it), np.percentile(it, 90), np.percentile(it, 99), it.max()))
h = w20[:, None] * tab[ix2] + w21[:, None] * tab[ix2 + 1]
seg = np.clip((h <= sie[:, None]).sum(axis=1) - 1, 0, 254)
cond = np.abs(sie) / np.abs(h[allm, seg + 1] - h[allm, seg])
for lo, hi in ((0, 1e2), (1e2, 1e3), (1e3, 1e4), (1e4, 3e4), (3e4, 1e5), (1e5, 1e9)):
    m = (cond >= lo) & (cond < hi)
    print("cond [%g,%g): frac %.3f  it mean %.1f p10 %d p90 %d max %d" % (lo, hi, m.mean(), it[m].mean(), np.percentile(it[m], 10), np.percentile(it[m], 90), it[m].max()))
# warp-level efficiency: mean/max over groups of 32 consecutive cells
w = it.reshape(-1, 32)
print("lane efficiency all cells: %.3f" % (w.mean() / w.max(axis=1).mean()))
hard = it[cond >= 2e4]; hard = hard[: hard.size // 32 * 32].reshape(-1, 32)
print("lane efficiency hard cells compacted: %.3f  (mean its %.1f, mean warp max %.1f)" % (hard.mean() / hard.max(axis=1).mean(), hard.mean(), hard.max(axis=1).mean()))
