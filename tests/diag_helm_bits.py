"""Diagnostic (not a test): bit-level agreement of Helmholtz FillEos and the analytic PTE
materials with the reference oracle.  usage: python tests/diag_helm_bits.py"""
import os, sys, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import numpy as np, torch
import __graft_entry__ as g
import oracle_ref
pkg = g.load_package(); L = pkg._lib; W = pkg.workloads; tq = pkg.thermalqs
ref = oracle_ref.load()
def dev(a): return torch.from_numpy(np.ascontiguousarray(a)).cuda()
def rep(name, got, want):
    got = np.asarray(got); want = np.asarray(want)
    same = (got.view(np.uint64) == want.view(np.uint64)) | (np.isnan(got) & np.isnan(want))
    with np.errstate(all="ignore"):
        rel = np.where(same, 0.0, np.abs(got - want) / np.maximum(np.abs(want), 1e-300))
    print(f"  {name}: bit-identical {100*same.mean():.4f}%  worst rel {np.nanmax(rel):.3e}  >1e-12: {(rel>1e-12).sum()}", flush=True)
COMP = [(1.0, 1.0), (4.0, 2.0), (12.0, 6.0), (16.0, 8.0), (56.0, 26.0)]
def lam(n, ncol, lT=-1.0):
    l = np.empty((n, ncol)); c = np.array(COMP)[np.arange(n) % 5]
    l[:, 0], l[:, 1] = c[:, 0], c[:, 1]
    if ncol > 2: l[:, 2] = lT
    return l
td = tempfile.mkdtemp()
table = pkg.tablegen.write_synthetic_helm_table(os.path.join(td, "helm.dat"))
n = 1 << 15
for coul in (True, False):
    print(f"Helmholtz FillEos(rho,T) coul={coul}")
    ge = pkg.Helmholtz(table, coul=coul)
    r = ref.Helmholtz(table, L.HelmOptions(1, 1, int(coul), 1, 1, 1, 1))
    rho_h, T_h = W.helm_inputs(n)
    out = tq.specific_internal_energy | tq.pressure | tq.specific_heat | tq.bulk_modulus
    lh = lam(n, 3)
    e_r, P_r, cv_r, b_r = (np.zeros(n) for _ in range(4))
    r.fill(rho_h.copy(), T_h.copy(), e_r, P_r, cv_r, b_r, out, lh, nthreads=0)
    ld = dev(lam(n, 3))
    e, P, cv, b = (torch.zeros(n, dtype=torch.float64, device="cuda") for _ in range(4))
    ge.FillEos(dev(rho_h), dev(T_h), e, P, cv, b, n, out, ld)
    torch.cuda.synchronize()
    for nm, a, w in (("e", e, e_r), ("P", P, P_r), ("cv", cv, cv_r), ("bmod", b, b_r)):
        rep(nm, a.cpu().numpy(), w)
    rep("lambda", ld.cpu().numpy().reshape(-1), lh.reshape(-1))
print("Helmholtz FillEos(rho,e) coul off, no lT slot")
ge = pkg.Helmholtz(table, coul=False, lambda_has_lT_slot=False)
r = ref.Helmholtz(table, L.HelmOptions(1, 1, 0, 1, 1, 1, 0))
rho_h, T_h = W.helm_inputs(n)
out = tq.specific_internal_energy | tq.pressure | tq.specific_heat | tq.bulk_modulus
e_h, P_r, cv_r, b_r = (np.zeros(n) for _ in range(4))
r.fill(rho_h.copy(), T_h.copy(), e_h, P_r, cv_r, b_r, out, lam(n, 3), nthreads=0)
out = tq.temperature | tq.pressure | tq.specific_heat | tq.bulk_modulus
T_r, P_r, cv_r, b_r = (np.zeros(n) for _ in range(4))
r.fill(rho_h.copy(), T_r, e_h.copy(), P_r, cv_r, b_r, out, lam(n, 2), az_only=True, nthreads=0)
T2, P2, cv2, b2 = (torch.zeros(n, dtype=torch.float64, device="cuda") for _ in range(4))
ge.FillEos(dev(rho_h), T2, dev(e_h), P2, cv2, b2, n, out, dev(lam(n, 2)))
torch.cuda.synchronize()
for nm, a, w in (("T", T2, T_r), ("P", P2, P_r), ("cv", cv2, cv_r), ("bmod", b2, b_r)):
    rep(nm, a.cpu().numpy(), w)
print("analytic PTE materials")
g3, r3 = pkg.closure.set_eos_3mat_analytic(), ref.set_eos_3mat()
names = ["Gruneisen", "DavisReactants", "DavisProducts"]
rho0 = [8.93, 1.89, 2.5]
for m in range(3):
    rho = rho0[m] * (0.5 + 1.5 * pkg.rng.uniform_np(11 + m, 0, n))
    T = 100.0 + 4000.0 * pkg.rng.uniform_np(21 + m, 0, n)
    e_r = r3[m].eval(L.E_FROM_RHO_T, rho, T, nthreads=0)
    o = torch.empty(n, dtype=torch.float64, device="cuda")
    for meth in range(10):
        in2 = T if meth in (0, 1, 4, 6, 8) else e_r
        want = r3[m].eval(meth, rho, in2, nthreads=0)
        g3[m]._eval(meth, dev(rho), dev(in2), o, n, None)
        torch.cuda.synchronize()
        rep(f"{names[m]} {L.METHOD_NAMES[meth]}", o.cpu().numpy(), want)
