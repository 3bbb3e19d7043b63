"""Diagnostic (not a test): where do GPU and reference PTE solves differ?"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import numpy as np, torch
import __graft_entry__ as g, oracle_ref
from helpers import dev
pkg = g.load_package(); L = pkg._lib; W = pkg.workloads; ref = oracle_ref.load()
g3, r3 = pkg.closure.set_eos_3mat_analytic(), ref.set_eos_3mat()
sie0 = [r3[m].eval(L.E_FROM_RHO_T, [W.PTE_RHO[m]], [W.PTE_T])[0] for m in range(3)]
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
solver = int(sys.argv[2]) if len(sys.argv) > 2 else 0
s_h = W.pte_state_np(n, sie0)
s_in = [a.copy() for a in s_h]
s_d = [dev(a) for a in s_h]
prm = pkg.closure.test_pte_params() if (len(sys.argv) <= 3 or sys.argv[3] == "test") else pkg.closure.MixParams()
st_r, res_r, tg = ref.pte_solve(solver, r3, *s_h, prm, nthreads=0)
status = torch.zeros(n, dtype=torch.int64, device="cuda")
resid = torch.zeros(n, dtype=torch.float64, device="cuda")
pkg.closure.pte_solve(solver, g3, *s_d, params=prm, status=status, residual=resid)
torch.cuda.synchronize()
st_g = status.cpu().numpy().view(np.uint64)
dg, dr = pkg.closure.decode_status(st_g), pkg.closure.decode_status(st_r)
for k in dg:
    print(k, "mismatch:", int((dg[k] != dr[k]).sum()), "/", n)
bad = np.nonzero((dg["converged"] != dr["converged"]) | (dg["slow_convergence_detected"] != dr["slow_convergence_detected"]))[0]
print("bad cells", bad[:10])
for i in bad[:5]:
    print(i, "in vfrac", s_in[1][i], "ref:", {k: int(dr[k][i]) for k in dr}, "gpu:", {k: int(dg[k][i]) for k in dg})
    print("  ref state", [a[i] for a in s_h], res_r[i])
    print("  gpu state", [a[i].cpu().numpy() for a in s_d], resid[i].item())
conv = dr["converged"] & dg["converged"]
for a, b, nm in zip(s_d, s_h, ("rho", "vfrac", "sie", "temp", "press")):
    e = np.abs(a.cpu().numpy()[conv] - b[conv]) / np.abs(b[conv])
    print(nm, "max rel err on converged:", e.max())
print("niter hist ref", np.bincount(dr["niter"].astype(int)))
print("niter hist gpu", np.bincount(dg["niter"].astype(int)))
