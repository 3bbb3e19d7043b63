"""Diagnostic (not a test): bit-level agreement of the PTE solvers with the reference oracle.
usage: python tests/diag_pte_bits.py [log2 cells]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import numpy as np, torch
import __graft_entry__ as g
import oracle_ref
pkg = g.load_package(); L = pkg._lib; W = pkg.workloads
ref = oracle_ref.load()
lg = int(sys.argv[1]) if len(sys.argv) > 1 else 13
n = 1 << lg
g3, r3 = pkg.closure.set_eos_3mat_analytic(), ref.set_eos_3mat()
sie0 = [r3[m].eval(L.E_FROM_RHO_T, [W.PTE_RHO[m]], [W.PTE_T])[0] for m in range(3)]
vf = W.pte_vfracs(n)
def dev(a): return torch.from_numpy(np.ascontiguousarray(a)).cuda()
for sname, solver in (("RhoT", L.PTE_RHOT), ("RhoU", L.PTE_RHOU), ("PT", L.PTE_PT)):
    for pname, params in (("test", pkg.closure.test_pte_params()), ("default", pkg.closure.MixParams())):
        nn = n if solver != L.PTE_PT else min(n, 1 << 11)
        rho_h, _, sie_h, temp_h, press_h = W.pte_state_np(nn, sie0)
        vfr = vf[:nn].copy()
        state_d = [dev(a) for a in (rho_h, vf[:nn].copy(), sie_h, temp_h, press_h)]
        st_r, res_r, _ = ref.pte_solve(solver, r3, rho_h, vfr, sie_h, temp_h, press_h, params, nthreads=0)
        status = torch.zeros(nn, dtype=torch.int64, device="cuda")
        resid = torch.zeros(nn, dtype=torch.float64, device="cuda")
        pkg.closure.pte_solve(solver, g3, *state_d, params=params, status=status, residual=resid)
        torch.cuda.synchronize()
        dg = pkg.closure.decode_status(status.cpu().numpy().view(np.uint64))
        dr = pkg.closure.decode_status(st_r)
        line = f"{sname}/{pname}: n={nn}"
        for k in ("converged", "slow_convergence_detected", "niter", "max_line_niter"):
            line += f" {k}={100*(dg[k]==dr[k]).mean():.3f}%"
        bits = []
        worst = []
        for got, want in zip(state_d, (rho_h, vfr, sie_h, temp_h, press_h)):
            gt = got.cpu().numpy()
            bits.append(100 * (gt.view(np.uint64) == want.view(np.uint64)).mean())
            worst.append(np.max(np.abs(gt - want) / np.maximum(np.abs(want), 1e-300)))
        line += " | bit-identical % (rho,vfrac,sie,T,P): " + " ".join(f"{b:.2f}" for b in bits)
        line += " | worst rel: " + " ".join(f"{w:.1e}" for w in worst)
        rb = 100 * (resid.cpu().numpy().view(np.uint64) == np.asarray(res_r).view(np.uint64)).mean()
        line += f" | residual bits {rb:.2f}%"
        print(line, flush=True)
