import os, sys, numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import __graft_entry__ as g
import oracle_ref
pkg = g.load_package(); L = pkg._lib; tg = pkg.tablegen
ref = oracle_ref.load()
raw = tg.vinet_sp5_tables()
r = ref.SpinerRhoT(raw)
n = 1 << 18
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
r0, r1 = tab[ix2], tab[ix2 + 1]          # [n,256]
h = w20[:, None] * r0 + w21[:, None] * r1
mono = (np.diff(h, axis=1) > 0).all(axis=1)
print("monotone rows:", mono.mean())
# segment: last i with h[i] <= sie
seg = np.clip((h <= sie[:, None]).sum(axis=1) - 1, 0, 254)
ar = np.arange(n)
h0, h1 = h[ar, seg], h[ar, seg + 1]
w11 = (sie - h0) / (h1 - h0)
dx = gT.dx[0]
lT = (seg * dx + gT.min[0]) + w11 * dx
T_dir = tg.fast_pow10(lT) - raw.struct.T_offset
rel = np.abs(T_dir - T_ref) / np.abs(T_ref)
dl = np.abs(lT - lT_ref)
cond = np.abs(sie) / np.abs(h1 - h0)
print("rel T err: max %.3e  frac>1e-12 %.4f  frac>1e-13 %.4f" % (rel.max(), (rel > 1e-12).mean(), (rel > 1e-13).mean()))
for c in (1e2, 1e3, 3e3, 1e4, 3e4, 1e5):
    m = cond < c
    print("cond<%g: frac %.4f  max rel %.3e  frac>1e-12 %.5f" % (c, m.mean(), rel[m].max(), (rel[m] > 1e-12).mean()))
print("T rel err vs true T input:", np.abs(T_ref - T).max() / 1, np.median(np.abs(T_ref - T) / T), np.median(np.abs(T_dir - T) / T))
np.savez("/tmp/proto/d.npz", rel=rel, cond=cond, dl=dl, rho=rho, T=T)
