#!/usr/bin/env python
"""Generates tests/golden/*.npz: outputs of the UNMODIFIED REFERENCE (oracle/_ref/libsgref.so,
built from /root/reference by oracle/Makefile) on the seeded synthetic inputs of
singularity-eos_b200/workloads.py.  Inputs are regenerated from their seeds by the tests, so
only outputs are stored.  TEST INFRASTRUCTURE; run where the reference is mounted:

    python tests/golden/make_golden.py
"""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

N_LOOKUP = 4096
N_PTE = 1024
N_HELM = 1024


def lookup_inputs(pkg, n=N_LOOKUP):
    return pkg.workloads.spiner_inputs(n, seeds=(pkg.workloads.SEED_ROOT_RHO,
                                                 pkg.workloads.SEED_ROOT_T))


def helm_inputs(pkg, n=N_HELM):
    rho, T = pkg.workloads.helm_inputs(n)
    lam = np.empty((n, 3))
    comp = np.array([(1., 1.), (4., 2.), (12., 6.), (16., 8.), (56., 26.)])[np.arange(n) % 5]
    lam[:, 0], lam[:, 1], lam[:, 2] = comp[:, 0], comp[:, 1], -1.0
    return rho, T, lam


def main():
    import __graft_entry__ as g
    import oracle_ref

    pkg = g.load_package()
    ref = oracle_ref.load()
    assert ref is not None, "build oracle/_ref first (make -C oracle)"
    L, W, tq = pkg._lib, pkg.workloads, pkg.thermalqs
    out = {}

    # config 1: IdealGas
    ig = ref.IdealGas(0.4, 7.16e6)
    rho, T = W.ideal_inputs(N_LOOKUP)
    out["ideal_P"] = ig.eval(L.P_FROM_RHO_T, rho, T)
    out["ideal_e"] = ig.eval(L.E_FROM_RHO_T, rho, T)

    # configs 2, 3a, 3b: Spiner tables from the analytic Vinet
    raw = pkg.tablegen.vinet_sp5_tables()
    rt, rs = ref.SpinerRhoT(raw), ref.SpinerRhoSie(raw)
    rho, T = lookup_inputs(pkg)
    lam = np.full((N_LOOKUP, 2), np.nan)
    out["rhot_P"] = rt.eval(L.P_FROM_RHO_T, rho, T, lam)
    out["rhot_lambda"] = lam.copy()
    sie = rt.eval(L.E_FROM_RHO_T, rho, T)
    out["rhot_e"] = sie
    lam = np.full((N_LOOKUP, 2), np.nan)
    out["rhot_T_of_e_cold"] = rt.eval(L.T_FROM_RHO_E, rho, sie, lam)
    out["rhot_T_of_e_lambda"] = lam.copy()
    lam1 = np.full((N_LOOKUP, 1), np.nan)
    out["rhosie_T_of_e"] = rs.eval(L.T_FROM_RHO_E, rho, sie, lam1)
    out["rhosie_P_of_e"] = rs.eval(L.P_FROM_RHO_E, rho, sie, lam1)

    # config 4: Helmholtz on the synthetic table, FillEos(rho,T) -> e,P,cv,bmod
    with tempfile.TemporaryDirectory() as d:
        path = pkg.tablegen.write_synthetic_helm_table(os.path.join(d, "helm.dat"))
        h = ref.Helmholtz(path, L.HelmOptions(1, 1, 1, 1, 1, 1, 1))
        rho, T, lam = helm_inputs(pkg)
        e, P, cv, b = (np.zeros(N_HELM) for _ in range(4))
        mask = tq.specific_internal_energy | tq.pressure | tq.specific_heat | tq.bulk_modulus
        h.fill(rho.copy(), T.copy(), e, P, cv, b, mask, lam)
        out["helm_e"], out["helm_P"], out["helm_cv"], out["helm_bmod"] = e, P, cv, b
        out["helm_lambda"] = lam.copy()

    # config 5: PTESolverRhoT / PTESolverPT, 3-material analytic mixture, default MixParams
    r3 = ref.set_eos_3mat()
    sie0 = [r3[m].eval(L.E_FROM_RHO_T, [W.PTE_RHO[m]], [W.PTE_T])[0] for m in range(3)]
    out["pte_sie0"] = np.array(sie0)
    for name, solver in (("rhot", L.PTE_RHOT), ("pt", L.PTE_PT)):
        st = [a.copy() for a in W.pte_state_np(N_PTE, sie0)]
        status, resid, _ = ref.pte_solve(solver, r3, *st, pkg.closure.MixParams(), nthreads=0)
        for a, nm in zip(st, ("rho", "vfrac", "sie", "temp", "press")):
            out[f"pte_{name}_{nm}"] = a
        out[f"pte_{name}_status"] = status
        out[f"pte_{name}_residual"] = resid
    np.savez_compressed(os.path.join(HERE, "reference_outputs.npz"), **out)
    print("wrote", os.path.join(HERE, "reference_outputs.npz"),
          os.path.getsize(os.path.join(HERE, "reference_outputs.npz")), "bytes")


if __name__ == "__main__":
    main()
