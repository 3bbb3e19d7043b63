"""smoke(): one small invocation of the hot path on cuda:0, checked against the oracle."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def run():
    import torch

    import __graft_entry__ as g
    import oracle_ref
    from helpers import assert_bits, assert_rel, dev

    assert torch.cuda.is_available(), "smoke() needs a CUDA device"
    torch.cuda.set_device(0)
    pkg = g.load_package()
    L = pkg._lib
    lib = L.load()
    ref = oracle_ref.load()
    assert ref is not None, "oracle/_ref/libsgref.so missing (build() compiles it)"
    before = lib.sgx_launch_count()

    # config 1: IdealGas vector lookups, bit-exact
    n = 1 << 16
    rho_h, T_h = pkg.workloads.ideal_inputs(n)
    ge, re_ = pkg.IdealGas(0.4, 7.16e6), ref.IdealGas(0.4, 7.16e6)
    rho_d, T_d = dev(rho_h), dev(T_h)
    out = torch.empty_like(rho_d)
    ge.PressureFromDensityTemperature(rho_d, T_d, out, n)
    assert_bits(out.cpu().numpy(), re_.eval(L.P_FROM_RHO_T, rho_h, T_h), "IdealGas P")

    # configs 2/3a: Spiner (rho,T) lookup and regula-falsi T(rho,e)
    raw = pkg.tablegen.vinet_sp5_tables()
    gs, rs = pkg.SpinerEOSDependsRhoT(raw), ref.SpinerRhoT(raw)
    rho_h, T_h = pkg.workloads.spiner_inputs(n)
    rho_d, T_d = dev(rho_h), dev(T_h)
    lam_d = dev(np.full((n, 2), np.nan))
    gs.InternalEnergyFromDensityTemperature(rho_d, T_d, out, n, lam_d)
    lam_h = np.full((n, 2), np.nan)
    e_h = rs.eval(L.E_FROM_RHO_T, rho_h, T_h, lam_h)
    assert_rel(out.cpu().numpy(), e_h, 1e-12, "Spiner e(rho,T)")
    lam_d.fill_(float("nan"))
    gs.TemperatureFromDensityInternalEnergy(rho_d, dev(e_h), out, n, lam_d)
    assert_rel(out.cpu().numpy(), rs.eval(L.T_FROM_RHO_E, rho_h, e_h, np.full((n, 2), np.nan)),
               1e-12, "Spiner T(rho,e) root find")

    # config 5: PTE RhoT on 4096 cells: identical flags, 1e-10 states
    W = pkg.workloads
    g3, r3 = pkg.closure.set_eos_3mat_analytic(), ref.set_eos_3mat()
    sie0 = [r3[m].eval(L.E_FROM_RHO_T, [W.PTE_RHO[m]], [W.PTE_T])[0] for m in range(3)]
    m = 4096
    s_h = W.pte_state_np(m, sie0)
    s_d = [dev(a) for a in s_h]
    prm = pkg.closure.MixParams()  # library defaults: every flag is reproducible
    st_r, _, _ = ref.pte_solve(L.PTE_RHOT, r3, *s_h, prm, nthreads=0)
    status = torch.zeros(m, dtype=torch.int64, device="cuda")
    pkg.closure.PTESolverRhoT(g3, *s_d, params=prm, status=status)
    torch.cuda.synchronize()
    st_g = status.cpu().numpy().view(np.uint64)
    assert ((st_g & 3) == (st_r & 3)).all(), "PTE convergence flags differ from the reference"
    conv = (st_r & 1) != 0
    for a, b, nm in zip(s_d, s_h, ("rho", "vfrac", "sie", "temp", "press")):
        assert_rel(a.cpu().numpy()[conv], b[conv], 1e-10, "PTE " + nm)
    assert (st_g == st_r).all(), "PTE SolverStatus words differ from the reference"

    # config 3b: SpinerEOSDependsRhoSie direct T(rho, sie)
    gq, rq = pkg.SpinerEOSDependsRhoSie(raw), ref.SpinerRhoSie(raw)
    lam1 = dev(np.full((n, 1), np.nan))
    gq.TemperatureFromDensityInternalEnergy(rho_d, dev(e_h), out, n, lam1)
    assert_rel(out.cpu().numpy(), rq.eval(L.T_FROM_RHO_E, rho_h, e_h, np.full((n, 1), np.nan)),
               1e-12, "SpinerRhoSie T(rho,e)")

    # config 4: Helmholtz FillEos(rho, T) -> e, P, cv, bmod on the synthetic table
    import tempfile
    tq = pkg.thermalqs
    with tempfile.TemporaryDirectory() as d:
        hp = pkg.tablegen.write_synthetic_helm_table(os.path.join(d, "helm.dat"))
        gh = pkg.Helmholtz(hp)
        rh = ref.Helmholtz(hp, L.HelmOptions(1, 1, 1, 1, 1, 1, 1))
    nh = 4096
    hr, hT = pkg.workloads.helm_inputs(nh)
    lam3 = np.tile(np.array([4.0, 2.0, -1.0]), (nh, 1))
    mask = tq.specific_internal_energy | tq.pressure | tq.specific_heat | tq.bulk_modulus
    e_r, P_r, cv_r, b_r = (np.zeros(nh) for _ in range(4))
    rh.fill(hr.copy(), hT.copy(), e_r, P_r, cv_r, b_r, mask, lam3.copy(), nthreads=0)
    he, hP, hcv, hb = (torch.zeros(nh, dtype=torch.float64, device="cuda") for _ in range(4))
    gh.FillEos(dev(hr), dev(hT), he, hP, hcv, hb, nh, mask, dev(lam3))
    for got, want, nm in ((he, e_r, "e"), (hP, P_r, "P"), (hcv, cv_r, "cv"), (hb, b_r, "bmod")):
        assert_rel(got.cpu().numpy(), want, 1e-12, "Helmholtz FillEos " + nm)

    # PTESolverPT on 512 cells: identical status words, 1e-10 states
    mp = 512
    p_h = W.pte_state_np(mp, sie0)
    p_d = [dev(a) for a in p_h]
    st_r, _, _ = ref.pte_solve(L.PTE_PT, r3, *p_h, prm, nthreads=0)
    status = torch.zeros(mp, dtype=torch.int64, device="cuda")
    pkg.closure.PTESolverPT(g3, *p_d, params=prm, status=status)
    torch.cuda.synchronize()
    assert (status.cpu().numpy().view(np.uint64) == st_r).all(), "PTESolverPT status differs"
    cpt = (st_r & 1) != 0
    for a, b, nm in zip(p_d, p_h, ("rho", "vfrac", "sie", "temp", "press")):
        assert_rel(a.cpu().numpy()[cpt], b[cpt], 1e-10, "PTESolverPT " + nm)

    launched = lib.sgx_launch_count() - before
    assert launched >= 9
    print(f"smoke ok: {launched} sgx kernels launched on {torch.cuda.get_device_name(0)}; "
          f"PTE converged {int(conv.sum())}/{m}")


if __name__ == "__main__":
    run()
