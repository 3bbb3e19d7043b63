"""Profiling driver (not a test): runs each hot kernel a few times at a size ncu can replay.
usage: python tests/prof_driver.py [lookup|root|pte|rhosie|ideal|all] [log2 cells]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import __graft_entry__ as g
pkg = g.load_package(); L = pkg._lib; W = pkg.workloads
what = sys.argv[1] if len(sys.argv) > 1 else "all"
lg = int(sys.argv[2]) if len(sys.argv) > 2 else 24
n = 1 << lg
dev = torch.device("cuda", 0)
if os.environ.get("PROF_INIT_NCCL") == "1":  # study: does a live NCCL communicator change kernel timings?
    import torch.distributed as dist
    torch.cuda.set_device(0)
    dist.init_process_group("nccl", init_method="tcp://127.0.0.1:29531", rank=0, world_size=1,
                            device_id=dev)
    t_ = torch.ones(1, device=dev); dist.all_reduce(t_); torch.cuda.synchronize()
    print("NCCL communicator up", flush=True)
raw = pkg.tablegen.vinet_sp5_tables()
rt = pkg.SpinerEOSDependsRhoT(raw)
rho, T = W.spiner_inputs(n, 0, xp=torch, device=dev)
out = torch.empty_like(rho); sie = torch.empty_like(rho)
lam = torch.full((n, 2), float("nan"), dtype=torch.float64, device=dev)
reps = 3
def timed(name, fn):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    print(f"{name}: {e0.elapsed_time(e1)/reps:.4f} ms for {n} cells", flush=True)
rt.InternalEnergyFromDensityTemperature(rho, T, sie, n, lam)
if what in ("lookup", "all"):
    timed("rhot_P_rhoT_lambda", lambda: rt.PressureFromDensityTemperature(rho, T, out, n, lam))
    timed("rhot_P_rhoT_nolambda", lambda: rt.PressureFromDensityTemperature(rho, T, out, n))
if what in ("root", "all"):
    def cold():
        lam.fill_(float("nan")); rt.TemperatureFromDensityInternalEnergy(rho, sie, out, n, lam)
    timed("rhot_T_rhoe_cold", cold)
    lam_w = lam.clone()  # holds the converged (lRho, lT) of the cold pass
    def warm():
        lam.copy_(lam_w); rt.TemperatureFromDensityInternalEnergy(rho, sie, out, n, lam)
    timed("rhot_T_rhoe_warm", warm)
if what in ("rhosie", "all"):
    rs = pkg.SpinerEOSDependsRhoSie(raw)
    lam1 = torch.full((n, 1), float("nan"), dtype=torch.float64, device=dev)
    timed("rhosie_T_rhoe", lambda: rs.TemperatureFromDensityInternalEnergy(rho, sie, out, n, lam1))
if what in ("ideal", "all"):
    ig = pkg.IdealGas(0.4, 7.16e6)
    timed("ideal_P", lambda: ig.PressureFromDensityTemperature(rho, T, out, n))
if what in ("helm", "all"):
    import tempfile
    tq = pkg.thermalqs
    hp = os.path.join(tempfile.gettempdir(), "sgx_helm_prof.dat")
    pkg.tablegen.write_synthetic_helm_table(hp)
    nh = 1 << (lg - 2)
    hr, hT = W.helm_inputs(nh, 0, xp=torch, device=dev)
    he, hP, hcv, hb = (torch.empty_like(hr) for _ in range(4))
    lam3 = torch.tensor([4.0, 2.0, -1.0], dtype=torch.float64, device=dev).repeat(nh, 1).contiguous()
    hm = pkg.Helmholtz(hp)
    mi = tq.specific_internal_energy | tq.pressure | tq.specific_heat | tq.bulk_modulus
    n_save = n; n = nh
    timed("helm_fill_rhoT", lambda: hm.FillEos(hr, hT, he, hP, hcv, hb, nh, mi, lam3))
    hm2 = pkg.Helmholtz(hp, coul=False, lambda_has_lT_slot=False)
    lam2 = lam3[:, :2].contiguous()
    hm2.FillEos(hr, hT, he, hP, hcv, hb, nh, mi, lam2)
    hT2 = torch.empty_like(hT)
    mii = tq.temperature | tq.pressure | tq.specific_heat | tq.bulk_modulus
    timed("helm_fill_rhoe_newton", lambda: hm2.FillEos(hr, hT2, he, hP, hcv, hb, nh, mii, lam2))
    n = n_save
if what in ("pte", "all"):
    npte = 1 << (lg - 4)
    g3 = pkg.closure.set_eos_3mat_analytic()
    sie0 = []
    for m in range(3):
        o = torch.empty(1, dtype=torch.float64, device=dev)
        g3[m].InternalEnergyFromDensityTemperature(torch.tensor([W.PTE_RHO[m]], dtype=torch.float64, device=dev), torch.tensor([W.PTE_T], dtype=torch.float64, device=dev), o, 1)
        sie0.append(float(o.item()))
    s0 = W.pte_state_torch(npte, sie0, dev)
    st = [a.clone() for a in s0]
    prm = pkg.closure.test_pte_params() if os.environ.get("PTE_PARAMS", "test") == "test" else pkg.closure.MixParams()
    def pte():
        for a, b in zip(st, s0): a.copy_(b)
        pkg.closure.PTESolverRhoT(g3, *st, params=prm)
    n_save = n; n = npte
    timed("pte_rhot[" + os.environ.get("SGX_PTE_KERNEL", "sm") + "]", pte)
    n = n_save
if what in ("ptetab",):
    # tabulated mixture (Spiner Cu + IdealGas + Gruneisen) with per-material lambda caches
    npte = 1 << (lg - 4)
    g_t = [rt, pkg.IdealGas(0.4, 7.16e6), pkg.closure.set_eos_3mat_analytic()[0]]
    rho0 = torch.tensor([8.93, 1.0e-2, 8.93], dtype=torch.float64, device=dev)
    rho_t = (rho0.repeat(npte, 1) * (0.9 + 0.2 * pkg.rng.uniform_torch(57, 0, npte * 3, dev).reshape(npte, 3))).contiguous()
    T_t = torch.full((npte,), 600.0, dtype=torch.float64, device=dev)
    sie_t = torch.empty((npte, 3), dtype=torch.float64, device=dev)
    col = torch.empty(npte, dtype=torch.float64, device=dev)
    for m, em in enumerate(g_t):
        if m == 0:
            em.InternalEnergyFromDensityTemperature(rho_t[:, m].contiguous(), T_t, col, npte, torch.full((npte, 2), float("nan"), dtype=torch.float64, device=dev))
        else:
            em.InternalEnergyFromDensityTemperature(rho_t[:, m].contiguous(), T_t, col, npte)
        sie_t[:, m] = col
    t0_ = [rho_t, W.pte_vfracs(npte, start=0, xp=torch, device=dev, nmat=3), sie_t,
           torch.full((npte, 3), 600.0, dtype=torch.float64, device=dev), torch.zeros((npte, 3), dtype=torch.float64, device=dev)]
    ts_ = [a.clone() for a in t0_]
    lam_t = torch.full((npte, 3, 3), float("nan"), dtype=torch.float64, device=dev)
    def ptetab():
        for a, b in zip(ts_, t0_): a.copy_(b)
        pkg.closure.pte_solve(L.PTE_RHOT, g_t, *ts_, params=pkg.closure.MixParams(), lambdas=lam_t)
    n_save = n; n = npte
    timed("pte_rhot_tabulated[" + os.environ.get("SGX_PTE_KERNEL", "sm") + "]", ptetab)
    n = n_save
