#!/usr/bin/env python
"""bench.py -- BASELINE.json's metric on B200: Gcells/s for the EOS lookup configurations and
PTE cell-solves/s, next to the reference's own CPU implementation timed on the host cores.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    torchrun --nproc-per-node N ... bench.py --gpus N ...     (one rank per GPU, NCCL)

One JSON line on stdout (rank 0).  A "step" is one pass of the headline hot path over one
batch: SpinerEOSDependsRhoT::PressureFromDensityTemperature over 2^26 cells per GPU
(BASELINE.json configs[1]); cells are sharded contiguously over ranks, tables replicated,
no data-path collective (weak scaling).  The other configurations are measured in the same
run and reported under "configs".
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "Gcells/s EOS lookups (SpinerEOSDependsRhoT P(rho,T), 2^26 cells/GPU); PTE cell solves/s in configs"
UNIT = "Gcells/s"
N_LOOKUP = 1 << 26
N_IDEAL = 1 << 20


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cells", type=int, default=N_LOOKUP, help="cells per GPU for lookup configs")
    ap.add_argument("--pte-cells", type=int, default=1 << 27,
                    help="PTE cells per GPU (BASELINE.json configs[4]: 2^27)")
    ap.add_argument("--pte-full", action="store_true", help="(default now) PTE at 2^27 cells")
    ap.add_argument("--quick", action="store_true", help="headline config only")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------
class ClockSampler:
    """SM clock + clock-event (throttle) reasons sampled DURING the timed region through NVML
    (nvidia_ml_py, ~2 ms period: the headline region is only K x 0.5 ms long); summary() cuts
    the stream into time windows.  Same quantities as the nvidia-smi clocks line of
    B200_PROFILING.md."""

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.rows = []  # (t, sm_mhz, reasons bitmask)
        self.max_mhz = None
        self._stop = threading.Event()
        self._t = None
        self.nv = None

    def _run(self):
        nv = self.nv
        h = nv.nvmlDeviceGetHandleByIndex(self.idx)
        try:
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
        except Exception:
            pass
        get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
            getattr(nv, "nvmlDeviceGetCurrentClocksThrottleReasons")
        while not self._stop.is_set():
            try:
                self.rows.append((time.perf_counter(),
                                  float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)),
                                  int(get_reasons(h))))
            except Exception:
                pass
            self._stop.wait(0.002)

    def start(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            self.nv = nv
        except Exception:
            return
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()

    def stop(self):
        self._stop.set()
        if self._t:
            self._t.join(timeout=5)

    def summary(self, t0=None, t1=None):
        # NVML clocks-event-reason bits (nvml.h)
        bits = {"sw_power_cap": 0x4, "hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20,
                "hw_thermal_slowdown": 0x40}
        sm, reasons = [], set()
        for t, mhz, mask in self.rows:
            if (t0 is not None and t < t0) or (t1 is not None and t > t1):
                continue
            sm.append(mhz)
            for name, bit in bits.items():
                if mask & bit:
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(reasons), "samples": len(sm)}


def pin_to_gpu_numa(gpu_index):
    """One process per GPU: run this rank on the CPUs NVML reports as local to its GPU, so the
    pinned host buffers of the e2e leg are first-touched on that NUMA node and the staging
    copies do not cross the socket interconnect.  Returns the CPU list, or None."""
    try:
        import pynvml as nv
        nv.nvmlInit()
        h = nv.nvmlDeviceGetHandleByIndex(gpu_index)
        ncpu = os.cpu_count() or 1
        words = nv.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        cpus = [64 * w + b for w, word in enumerate(words) for b in range(64)
                if (int(word) >> b) & 1]
        allowed = os.sched_getaffinity(0)
        cpus = [c for c in cpus if c in allowed]
        if cpus:
            os.sched_setaffinity(0, cpus)
            return cpus
    except Exception:
        pass
    return None


def headline_traffic(cells):
    """DRAM bytes per launch of the headline kernel from the committed `ncu --set full`
    capture (profiles/headline_traffic.json), if it was taken at this launch size."""
    try:
        with open(os.path.join(ROOT, "profiles", "headline_traffic.json")) as fh:
            t = json.load(fh)
        if int(t["cells"]) == int(cells):
            return float(t["dram_bytes_read"]) + float(t["dram_bytes_write"])
    except Exception:
        pass
    return None


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


# ------------------------------------------------------------------------------------------
def time_steps(torch, fn, steps, warmup, dist=None):
    """W warm-ups, then K steps between CUDA events on the launching (current) stream,
    bracketed by barrier + synchronize; returns max-over-ranks seconds for the K steps."""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True)
    e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    secs = e0.elapsed_time(e1) * 1e-3
    if dist is not None:
        t = torch.tensor([secs], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        secs = float(t.item())
    return secs


def run_ours(args):
    import torch

    import __graft_entry__ as g

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py (impl=ours) needs a CUDA device: there is no CPU path")
    torch.cuda.set_device(local)
    all_cpus = sorted(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else []
    numa = pin_to_gpu_numa(local) if world > 1 else None
    dist = None
    if world > 1:
        import torch.distributed as dist_mod

        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"  # keep NCCL's version banner off stdout
        dist_mod.init_process_group("nccl", device_id=torch.device("cuda", local))
        dist = dist_mod
    pkg = g.load_package()
    L = pkg._lib
    lib = L.load()
    W = pkg.workloads
    dev = torch.device("cuda", local)
    K, Wu = args.steps, max(args.warmup, 3)
    n = args.cells
    start = rank * n  # contiguous shard of the global index space
    peak, peak_src = measured_peak_gbs()
    results = {}

    raw = pkg.tablegen.vinet_sp5_tables()
    eos_rt = pkg.SpinerEOSDependsRhoT(raw)
    rho, T = W.spiner_inputs(n, start, xp=torch, device=dev)
    out = torch.empty_like(rho)
    lam = torch.full((n, 2), float("nan"), dtype=torch.float64, device=dev)

    # ---- headline: config 2, P(rho,T) with (lRho,lT) lambda write-back: 40 B/cell ------
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.1)
    launches0 = lib.sgx_launch_count()
    t_head0 = time.perf_counter()
    secs = time_steps(torch, lambda: eos_rt.PressureFromDensityTemperature(rho, T, out, n, lam),
                      K, Wu, dist)
    t_head1 = time.perf_counter()
    launches = lib.sgx_launch_count() - launches0 - Wu
    cells_total = n * world * K
    value = cells_total / secs / 1e9
    bytes_per_cell = 40
    achieved = n * bytes_per_cell / (secs / K) / 1e9
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": headline_traffic(n),
                "algorithmic_bytes_per_launch": n * bytes_per_cell, "peak_source": peak_src,
                "kernel": "eval_kernel<P_FROM_RHO_T, SpinerRhoT, VEC>",
                "algorithmic_bytes_per_cell": bytes_per_cell}
    results["spiner_rhot_P_rhoT_lambda"] = {"Gcells_per_s": value, "GBps_per_gpu": achieved,
                                            "frac_hbm": achieved / peak, "cells_per_gpu": n}

    def lookup(name, fn, bpc, cells=n, steps=K):
        s = time_steps(torch, fn, steps, Wu, dist)
        gbs = cells * bpc / (s / steps) / 1e9
        results[name] = {"Gcells_per_s": cells * world * steps / s / 1e9, "GBps_per_gpu": gbs,
                         "frac_hbm": gbs / peak, "cells_per_gpu": cells,
                         "algorithmic_bytes_per_cell": bpc}

    if not args.quick:
        lookup("spiner_rhot_P_rhoT_nolambda",
               lambda: eos_rt.PressureFromDensityTemperature(rho, T, out, n), 24)
        # config 3a: regula-falsi T(rho,e), cold start (NaN lambda) and warm start
        sie = torch.empty_like(rho)
        eos_rt.InternalEnergyFromDensityTemperature(rho, T, sie, n, lam)
        lam_warm = lam.clone()

        def cold():
            lam.fill_(float("nan"))
            eos_rt.TemperatureFromDensityInternalEnergy(rho, sie, out, n, lam)

        def warm():
            lam.copy_(lam_warm)
            eos_rt.TemperatureFromDensityInternalEnergy(rho, sie, out, n, lam)
        ks = max(3, K // 4)
        eos_rt.root_stats(reset=True)
        lookup("spiner_rhot_T_rhoe_rootfind_cold(+lambda refill)", cold, 40, steps=ks)
        seen, listed = eos_rt.root_stats(reset=True)
        results["spiner_rhot_T_rhoe_rootfind_cold(+lambda refill)"].update({
            "root_mode": os.environ.get("SGX_ROOT_MODE", "hybrid"),
            "cells_through_reference_iterate_frac": listed / max(1, seen)})
        lookup("spiner_rhot_T_rhoe_rootfind_warm(+lambda refill)", warm, 40, steps=ks)
        # the reference iterate for every cell (SGX_ROOT_MODE=exact), same inputs
        os.environ["SGX_ROOT_MODE"] = "exact"
        try:
            lookup("spiner_rhot_T_rhoe_rootfind_cold_exact_iterate", cold, 40, steps=3)
        finally:
            os.environ.pop("SGX_ROOT_MODE", None)
        # config 3b: RhoSie direct lookup
        eos_rs = pkg.SpinerEOSDependsRhoSie(raw)
        lam1 = torch.full((n, 1), float("nan"), dtype=torch.float64, device=dev)
        lookup("spiner_rhosie_T_rhoe", lambda: eos_rs.TemperatureFromDensityInternalEnergy(
            rho, sie, out, n, lam1), 32)
        # a modified handle, ScaledEOS(ShiftedEOS(SpinerEOSDependsRhoT)), through
        # the run-time-dispatched modifier kernel (csrc/sgx_modifiers.cu)
        eos_mod = pkg.ScaledEOS(pkg.ShiftedEOS(eos_rt, 2.0e8), 1.05)
        lookup("modified_scaled_shifted_spiner_rhot_P_rhoT_lambda",
               lambda: eos_mod.PressureFromDensityTemperature(rho, T, out, n, lam), 40, steps=5)
        del lam1, lam_warm, eos_mod
        # config 1: IdealGas at 2^20 (launch-latency bound; CUDA graph of 100 pairs) and 2^26
        ideal = pkg.IdealGas(0.4, 7.16e6)
        r1, t1 = W.ideal_inputs(N_IDEAL, rank * N_IDEAL, xp=torch, device=dev)
        p1, e1 = torch.empty_like(r1), torch.empty_like(r1)

        def pair_small():
            ideal.PressureFromDensityTemperature(r1, t1, p1, N_IDEAL)
            ideal.InternalEnergyFromDensityTemperature(r1, t1, e1, N_IDEAL)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            pair_small()
        torch.cuda.current_stream().wait_stream(side)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            for _ in range(100):
                pair_small()
        s = time_steps(torch, graph.replay, 10, Wu, dist)
        results["ideal_P+e_1M_cudagraph"] = {
            "Gcells_per_s": N_IDEAL * world * 1000 / s / 1e9, "us_per_pair": s / 1000 * 1e6,
            "note": "2^20 cells fit L2; launch-latency bound (SURVEY 8d)",
            "cells_per_gpu": N_IDEAL}

        def pair_big():
            ideal.PressureFromDensityTemperature(rho, T, out, n)
            ideal.InternalEnergyFromDensityTemperature(rho, T, sie, n)
        lookup("ideal_P_then_e_64M", pair_big, 48)
        lookup("ideal_fused_P+e_64M",
               lambda: ideal.PressureEnergyFromDensityTemperature(rho, T, out, sie, n), 32)
        del sie
    del lam

    # ---- config 4: Helmholtz FillEos on the synthetic table (the shipped one is absent) ----
    if not args.quick:
        import tempfile
        tq = pkg.thermalqs
        hpath = os.path.join(tempfile.gettempdir(), f"sgx_helm_synthetic_{os.getpid()}.dat")
        pkg.tablegen.write_synthetic_helm_table(hpath)
        nh = n  # BASELINE.json configs[3]: 2^26 cells
        h_rho, h_T = W.helm_inputs(nh, rank * nh, xp=torch, device=dev)
        h_e, h_P, h_cv, h_b = (torch.empty_like(h_rho) for _ in range(4))
        comp = torch.tensor([[1., 1.], [4., 2.], [12., 6.], [16., 8.], [56., 26.]],
                            dtype=torch.float64, device=dev)
        idx = torch.arange(nh, device=dev) % 5
        lam3 = torch.cat([comp[idx], torch.full((nh, 1), -1.0, dtype=torch.float64, device=dev)],
                         dim=1).contiguous()
        helm = pkg.Helmholtz(hpath)
        mode_i = tq.specific_internal_energy | tq.pressure | tq.specific_heat | tq.bulk_modulus
        kh = max(3, K // 4)
        lookup("helmholtz_FillEos_rhoT(e,P,cv,bmod)",
               lambda: helm.FillEos(h_rho, h_T, h_e, h_P, h_cv, h_b, nh, mode_i, lam3), 72,
               cells=nh, steps=kh)
        # mode (ii): (rho,e) -> T,P,cv,bmod, Newton on e(T); Coulomb off (the reference test's
        # variant: with Coulomb on e(T) is non-monotone on ~6 % of this box), no lT slot
        helm2 = pkg.Helmholtz(hpath, coul=False, lambda_has_lT_slot=False)
        helm2.FillEos(h_rho, h_T, h_e, h_P, h_cv, h_b, nh, mode_i, lam3[:, :2].contiguous())
        lam2 = lam3[:, :2].contiguous()
        h_T2 = torch.empty_like(h_T)
        mode_ii = tq.temperature | tq.pressure | tq.specific_heat | tq.bulk_modulus
        lookup("helmholtz_FillEos_rhoe(T,P,cv,bmod)_newton_coul_off",
               lambda: helm2.FillEos(h_rho, h_T2, h_e, h_P, h_cv, h_b, nh, mode_ii, lam2), 64,
               cells=nh, steps=kh)
        del h_rho, h_T, h_e, h_P, h_cv, h_b, lam3, lam2, h_T2, helm, helm2
        try:
            os.remove(hpath)
        except OSError:
            pass

    # ---- config 5: PTE RhoT -------------------------------------------------------------
    if not args.quick:
        npte = args.pte_cells
        g3 = pkg.closure.set_eos_3mat_analytic()
        sie0 = []
        for m in range(3):
            o = torch.empty(1, dtype=torch.float64, device=dev)
            g3[m].InternalEnergyFromDensityTemperature(
                torch.tensor([W.PTE_RHO[m]], dtype=torch.float64, device=dev),
                torch.tensor([W.PTE_T], dtype=torch.float64, device=dev), o, 1)
            sie0.append(float(o.item()))
        state0 = W.pte_state_torch(npte, sie0, dev, start=rank * npte)
        state = [a.clone() for a in state0]
        status = torch.zeros(npte, dtype=torch.int64, device=dev)
        stats = torch.zeros(L.PTE_NSTATS, dtype=torch.int64, device=dev)
        prm = pkg.closure.test_pte_params()

        def pte():
            for a, b in zip(state, state0):
                a.copy_(b)
            pkg.closure.PTESolverRhoT(g3, *state, params=prm, status=status, stats=stats)
        ks = 3
        s = time_steps(torch, pte, 1, 2, dist)  # warm-up passes (allocator pools, clocks)
        stats.zero_()
        s = time_steps(torch, pte, ks, 0, dist)
        if dist is not None:
            dist.all_reduce(stats)  # the optional convergence-statistics all-reduce (NCCL)
        st = stats.cpu().numpy()
        results["pte_rhot_3mat"] = {
            "cell_solves_per_s": npte * world * ks / s, "cells_per_gpu": npte,
            "mix_params": "test/test_pte.cpp:164-175 (tolerances 1e-16)",
            "converged_frac": float(st[1]) / max(1, float(st[0])),
            "niter_hist": [int(x) for x in st[8:24]],
            "algorithmic_bytes_per_cell": 224,
            "note": "includes the 5-array state reset copy each step (in-place solver)"}
        # secondary: library-default MixParams, and PTESolverPT, on a 2^22-cell slice
        nsec = min(npte, 1 << 22)
        sec0 = [a[:nsec] for a in state0]
        sec = [a[:nsec] for a in state]
        prm_def = pkg.closure.MixParams()

        def pte_def():
            for a, b in zip(sec, sec0):
                a.copy_(b)
            pkg.closure.PTESolverRhoT(g3, *sec, params=prm_def)
        s = time_steps(torch, pte_def, ks, 1, dist)
        results["pte_rhot_3mat_default_mixparams"] = {
            "cell_solves_per_s": nsec * world * ks / s, "cells_per_gpu": nsec}
        npt = min(npte, 1 << 24)  # PTESolverPT at 2^24 cells

        def pte_pt():
            for a, b in zip(state, state0):
                a[:npt].copy_(b[:npt])
            pkg.closure.PTESolverPT(g3, *[a[:npt] for a in state], params=prm_def)
        s = time_steps(torch, pte_pt, 2, 1, dist)
        results["pte_pt_3mat_default_mixparams"] = {
            "cell_solves_per_s": npt * world * 2 / s, "cells_per_gpu": npt}
        # the runtime-dispatched (Variant-like, DynEos) path: 2 and 4 materials, 2^22 cells
        for nm, mats in ((2, (0, 2)), (4, (0, 1, 2, 0))):
            ng = min(npte, 1 << 22)
            eos_g = [g3[m] for m in mats]
            vf = W.pte_vfracs(ng, start=rank * ng, xp=torch, device=dev, nmat=nm)
            rho_g = torch.tensor([W.PTE_RHO[m] for m in mats], dtype=torch.float64,
                                 device=dev).repeat(ng, 1).contiguous()
            sie_g = torch.tensor([sie0[m] for m in mats], dtype=torch.float64,
                                 device=dev).repeat(ng, 1).contiguous()
            tmp_g = torch.full((ng, nm), W.PTE_T, dtype=torch.float64, device=dev)
            prs_g = torch.zeros((ng, nm), dtype=torch.float64, device=dev)
            g0 = [rho_g, vf, sie_g, tmp_g, prs_g]
            gs = [a.clone() for a in g0]
            stg = torch.zeros(ng, dtype=torch.int64, device=dev)

            def pte_gen():
                for a, b in zip(gs, g0):
                    a.copy_(b)
                pkg.closure.PTESolverRhoT(eos_g, *gs, params=prm_def, status=stg)
            s = time_steps(torch, pte_gen, ks, 1, dist)
            results[f"pte_rhot_{nm}mat_generic_default_mixparams"] = {
                "cell_solves_per_s": ng * world * ks / s, "cells_per_gpu": ng,
                "converged_frac": float((stg & 1).double().mean().item()),
                "materials": [type(e).__name__ for e in eos_g]}
            del g0, gs, stg
        # a production-like mixed cell -- tabulated copper (SpinerEOSDependsRhoT,
        # per-material lambda caches) + IdealGas + Gruneisen, 2^22 cells
        nt = min(npte, 1 << 22)
        eos_t = [eos_rt, pkg.IdealGas(0.4, 7.16e6), g3[0]]
        rho0_t = torch.tensor([8.93, 1.0e-2, 8.93], dtype=torch.float64, device=dev)
        rho_t = (rho0_t.repeat(nt, 1) *
                 (0.9 + 0.2 * pkg.rng.uniform_torch(57, rank * nt * 3, nt * 3, dev)
                  .reshape(nt, 3))).contiguous()
        T_t = torch.full((nt,), 600.0, dtype=torch.float64, device=dev)
        sie_t = torch.empty((nt, 3), dtype=torch.float64, device=dev)
        col = torch.empty(nt, dtype=torch.float64, device=dev)
        lam_t = torch.full((nt, 3, 3), float("nan"), dtype=torch.float64, device=dev)
        for m, em in enumerate(eos_t):
            if m == 0:
                em.InternalEnergyFromDensityTemperature(rho_t[:, m].contiguous(), T_t, col, nt,
                                                        lam_t[:, 0, :2].contiguous())
            else:
                em.InternalEnergyFromDensityTemperature(rho_t[:, m].contiguous(), T_t, col, nt)
            sie_t[:, m] = col
        t0_ = [rho_t, W.pte_vfracs(nt, start=rank * nt, xp=torch, device=dev, nmat=3), sie_t,
               torch.full((nt, 3), 600.0, dtype=torch.float64, device=dev),
               torch.zeros((nt, 3), dtype=torch.float64, device=dev)]
        ts_ = [a.clone() for a in t0_]
        stt = torch.zeros(nt, dtype=torch.int64, device=dev)

        def pte_tab():
            for a, b in zip(ts_, t0_):
                a.copy_(b)
            lam_t.fill_(float("nan"))
            pkg.closure.pte_solve(L.PTE_RHOT, eos_t, *ts_, params=prm_def, status=stt,
                                  lambdas=lam_t)
        s = time_steps(torch, pte_tab, ks, 1, dist)
        results["pte_rhot_tabulated_3mat_default_mixparams"] = {
            "cell_solves_per_s": nt * world * ks / s, "cells_per_gpu": nt,
            "converged_frac": float((stt & 1).double().mean().item()),
            "materials": ["SpinerEOSDependsRhoT (Vinet Cu table, lambda cache)", "IdealGas",
                          "Gruneisen"],
            "kernel": "pte_rhot_sm_kernel<3, TabEosT<false>> (lane-refill state machine, per-lane caches)"}
        del t0_, ts_, stt, lam_t, sie_t, rho_t
        # the get_sg_eos batch driver, (rho, e) input, on HOST arrays in and out
        # (the reference's memory contract: it mirrors every array to the device and back)
        nsg = 1 << 20

        def e_gpu(m, rho_col, T_col):
            o = torch.empty(rho_col.size, dtype=torch.float64, device=dev)
            g3[m].InternalEnergyFromDensityTemperature(
                torch.from_numpy(rho_col).to(dev), torch.from_numpy(T_col).to(dev), o,
                rho_col.size)
            return o.cpu().numpy()
        a0, offs = sg_eos_inputs(pkg, nsg, e_gpu)
        t_sg = []
        for k in range(5):
            a1 = {kk: vv.copy(order="K") for kk, vv in a0.items()}
            if dist is not None:
                dist.barrier()
            t0 = time.perf_counter()
            pkg.closure.get_sg_eos(g3, [1, 2, 3], offs, **a1, input_int=1, mass_frac_cutoff=1e-8)
            dt = time.perf_counter() - t0
            if dist is not None:
                tt = torch.tensor([dt], dtype=torch.float64, device=dev)
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
                dt = float(tt.item())
            if k >= 2:
                t_sg.append(dt)
        results["get_sg_eos_rho_e_3mat_host"] = {
            "cells_per_s": nsg * world / float(np.median(t_sg)), "cells_per_gpu": nsg,
            "temp_eV_mean": float(a1["temp"].mean()),
            "api": "sgx_get_sg_eos: host arrays in and out, 1 launch of get_sg_eos_kernel<rho_e>; "
                   "2 warm-up + 3 timed calls, median, max over ranks"}
        del a0, a1
        # e2e for config 5: host buffers through sgx_pte_solve_host (chunked, 3 streams)
        npe = min(npte, 1 << 22)
        hst0 = [a[:npe].cpu() for a in state0]
        hst = [torch.empty_like(a).pin_memory() for a in hst0]
        hstatus = torch.zeros(npe, dtype=torch.int64).pin_memory()

        def pte_host():
            for a, b in zip(hst, hst0):
                a.copy_(b)
            pkg.closure.pte_solve_host(L.PTE_RHOT, g3, *[a.numpy() for a in hst], params=prm,
                                       status=hstatus.numpy().view(np.uint64))
        pte_host()
        pte_host()
        t_pass = []
        for _ in range(5):
            for a, b in zip(hst, hst0):
                a.copy_(b)
            if dist is not None:
                dist.barrier()
            t0 = time.perf_counter()
            pkg.closure.pte_solve_host(L.PTE_RHOT, g3, *[a.numpy() for a in hst], params=prm,
                                       status=hstatus.numpy().view(np.uint64))
            dt = time.perf_counter() - t0
            if dist is not None:
                tt = torch.tensor([dt], dtype=torch.float64, device=dev)
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
                dt = float(tt.item())
            t_pass.append(dt)
        results["pte_rhot_3mat_e2e_host"] = {
            "cell_solves_per_s": npe * world / float(np.median(t_pass)), "cells_per_gpu": npe,
            "passes": len(t_pass), "statistic": "median pass, max over ranks",
            "h2d_bytes_per_cell": 120, "d2h_bytes_per_cell": 128,
            "api": "sgx_pte_solve_host with pinned host arrays"}
        del state, state0, status, hst, hst0

    # ---- e2e: host buffers through the C ABI (_host entry), H2D + D2H inside the timing --
    # The headline workload itself: 2^26 cells per step per GPU, from the caller's host arrays.
    ne = n
    rho_h, T_h = W.spiner_inputs(ne, start)
    rho_p = torch.from_numpy(rho_h).pin_memory().numpy()
    T_p = torch.from_numpy(T_h).pin_memory().numpy()
    out_p = torch.empty(ne, dtype=torch.float64).pin_memory().numpy()
    lam_p = torch.full((ne, 2), float("nan"), dtype=torch.float64).pin_memory().numpy()

    def timed_host(fn, reps):
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        ts = []
        for _ in range(reps):
            if dist is not None:
                dist.barrier()
            t0 = time.perf_counter()
            fn()
            dt = time.perf_counter() - t0
            if dist is not None:
                tt = torch.tensor([dt], dtype=torch.float64, device=dev)
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
                dt = float(tt.item())
            ts.append(dt)
        return float(np.sum(ts))
    ke = 5
    t_e2e = timed_host(lambda: eos_rt.eval_host(L.P_FROM_RHO_T, rho_p, T_p, out_p, lam_p), ke)
    e2e = {"value": ne * world * ke / t_e2e / 1e9, "unit": UNIT,
           # rho, T in; P and the (lRho, lT) lambda rows out.  The lambda rows are write-only for
           # a (rho, T) lookup, so the library does not copy them to the device (lambda_io)
           "h2d_bytes_per_step": ne * (8 + 8), "d2h_bytes_per_step": ne * (8 + 16),
           "cells_per_step_per_gpu": ne,
           "api": "sgx_eos_eval_host(P_FROM_RHO_T) with pinned host arrays",
           "host_cpus_of_rank0": (f"{len(numa)} CPUs local to the GPU (NVML affinity)"
                                  if numa else "unrestricted")}
    if not args.quick:
        # what a host code with ordinary (pageable) arrays sees: the library page-locks the
        # caller's ranges for the duration of the call (cudaHostRegister, HostPin in
        # csrc/sgx_internal.hpp); SGX_HOST_PIN=none is read once per process, so only the
        # default is timed here
        del rho_p, T_p, out_p, lam_p
        out_g = np.empty(ne)
        lam_g = np.full((ne, 2), np.nan)
        t_pg = timed_host(lambda: eos_rt.eval_host(L.P_FROM_RHO_T, rho_h, T_h, out_g, lam_g), 3)
        e2e["pageable_host_arrays_Gcells_per_s"] = ne * world * 3 / t_pg / 1e9
        e2e["pageable_note"] = ("numpy (pageable) arrays; cudaHostRegister + unregister of all "
                                "four ranges inside every timed call")
        del out_g, lam_g

    clocks = None
    if rank == 0:
        sampler.stop()
        clocks = sampler.summary(t_head0, t_head1)  # warm-up + timed region of the headline
        clocks["all_configs"] = sampler.summary()
    # FP64 pipe peak of this GPU, for the FP64-bound configs (3a, 4, 5); pipe utilisation of
    # those kernels comes from the committed ncu captures (profiles/r1h_*_details.txt)
    import ctypes
    fp64 = {}
    for name, fused in (("dfma_tflops", 1), ("dmul_dadd_tflops", 0)):
        v = ctypes.c_double(0.0)
        if lib.sgx_microbench_fp64(fused, ctypes.byref(v)) == 0:
            fp64[name] = v.value
    try:  # FP64-pipe utilisation: only what a committed ncu capture says, named with its file
        with open(os.path.join(ROOT, "profiles", "fp64_pipe_from_ncu.json")) as fh:
            fp64["ncu_pipe_fp64_pct"] = json.load(fh)
    except Exception:
        pass
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K,
        "warmup": Wu, "ms_per_step": secs / K * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "SpinerEOSDependsRhoT PressureFromDensityTemperature, 256x256 "
                               "NQT-log table from analytic Vinet, lambda (lRho,lT) write-back",
                   "cells_per_gpu": n, "global_cells": n * world,
                   "l2_policy": "inputs (1 GiB/GPU) larger than L2; no flush needed",
                   "parallelism": f"cells sharded contiguously over {world} GPU(s), tables replicated"},
        "roofline": roofline, "e2e": e2e, "gpu_launches": int(launches),
        "clocks": clocks, "fp64_peak": fp64, "configs": results,
    }
    if all_cpus:
        os.sched_setaffinity(0, all_cpus)  # the CPU baseline uses every host core
    if rank == 0 and not args.no_cpu:
        line["cpu_baseline"] = cpu_baseline(pkg, raw, sample=1 << 24, reps=5)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------
def host_threads():
    """Threads for the CPU legs: every core this process may run on (torchrun exports
    OMP_NUM_THREADS=1, which would otherwise cap the OpenMP baseline at one thread)."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def baseline_ref():
    """The library the CPU legs time: the Kokkos-OpenMP build of the reference harness when it
    exists (oracle/_ref/libsgref_kokkos.so), else the serial-portableFor + OpenMP-chunk build."""
    import oracle_ref

    ref = oracle_ref.load_kokkos()
    if ref is not None:
        return ref, ("reference vector API under PORTABILITY_STRATEGY_KOKKOS, vendored Kokkos 4.4 "
                     "OpenMP backend, g++ -O3 -DNDEBUG, OMP_PROC_BIND=%s OMP_PLACES=%s; PTE 3x3 "
                     "solves through Eigen" % (os.environ.get("OMP_PROC_BIND"),
                                               os.environ.get("OMP_PLACES")))
    ref = oracle_ref.load()
    return ref, ("OpenMP chunks over the reference vector API (PORTABILITY_STRATEGY_NONE), "
                 "g++ -O2; PTE 3x3 solves through Eigen")


def sg_eos_inputs(pkg, ncell, e_of):
    """Inputs of get_sg_eos (rho, e) for the test/test_pte.cpp mixture (Gruneisen / DavisReactants /
    DavisProducts): ncell cells, three materials each, cell_dim = ncell, identity offsets.
    e_of(m, rho, T) -> specific internal energies of material m (numpy)."""
    W = pkg.workloads
    vf = W.pte_vfracs(ncell)
    rho = np.array(W.PTE_RHO)[None, :] * (0.95 + 0.1 * pkg.rng.uniform_np(91, 0, ncell * 3)
                                           .reshape(ncell, 3))
    T0 = np.full(ncell, float(W.PTE_T))
    e_m = np.stack([e_of(m, rho[:, m].copy(), T0) for m in range(3)], axis=1)
    mass = rho * vf
    msum = mass.sum(axis=1)
    a = dict(press=np.zeros(ncell), pmax=np.full(ncell, -1e300), vol=np.ones(ncell),
             spvol=1.0 / msum, sie=(mass * e_m).sum(axis=1) / msum, temp=np.zeros(ncell),
             bmod=np.zeros(ncell), dpde=np.zeros(ncell), cv=np.zeros(ncell),
             frac_mass=np.asfortranarray(mass / msum[:, None]),
             frac_vol=np.zeros((ncell, 3), order="F"), frac_ie=np.zeros((ncell, 3), order="F"))
    return a, (np.arange(ncell, dtype=np.int32) + 1)


def cpu_baseline(pkg, raw, sample, reps):
    """The reference's own code (oracle/_ref) on the host cores, bounded sample of the headline
    workload.  This is the checker being *timed*, never part of the product path."""
    cores = host_threads()
    os.environ["OMP_NUM_THREADS"] = str(cores)
    ref, how = baseline_ref()
    if ref is None:
        return {"value": None, "unit": UNIT, "cores": 0, "kind": "reference",
                "sample": "oracle/_ref/libsgref*.so not present"}
    L = pkg._lib
    eos = ref.SpinerRhoT(raw)
    rho, T = pkg.workloads.spiner_inputs(sample)
    lam = np.full((sample, 2), np.nan)
    for _ in range(2):
        eos.eval(L.P_FROM_RHO_T, rho, T, lam, nthreads=cores)
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        eos.eval(L.P_FROM_RHO_T, rho, T, lam, nthreads=cores)
        ts.append(time.perf_counter() - t0)
    out = {"value": sample / float(np.median(ts)) / 1e9, "unit": UNIT, "cores": cores,
           "kind": "reference",
           "sample": f"{sample} cells of the headline workload, 2 warm-up + {reps} timed calls, "
                     f"median; {how}"}
    # config 5 beside it: the reference PTESolverRhoT on 2^22 cells of the same
    # workload, 3 warm-up + up to 10 timed passes (stopping after ~60 s), median
    try:
        W = pkg.workloads
        r3 = ref.set_eos_3mat()
        sie0 = [r3[m].eval(L.E_FROM_RHO_T, [W.PTE_RHO[m]], [W.PTE_T], nthreads=1)[0]
                for m in range(3)]
        npte = 1 << 22
        st0 = W.pte_state_np(npte, sie0)
        prm = pkg.closure.test_pte_params()
        tp, t_begin = [], time.perf_counter()
        for k in range(13):
            st = [a.copy() for a in st0]
            t0 = time.perf_counter()
            ref.pte_solve(L.PTE_RHOT, r3, *st, prm, nthreads=cores)
            if k >= 3:
                tp.append(time.perf_counter() - t0)
            if time.perf_counter() - t_begin > 60.0 and len(tp) >= 3:
                break
        out["pte_rhot_cell_solves_per_s"] = npte / float(np.median(tp))
        out["pte_sample"] = (f"{npte} cells, 3 warm-up + {len(tp)} timed passes, median, "
                             f"{cores} threads")
    except Exception as exc:  # the headline baseline stands on its own
        out["pte_error"] = repr(exc)
    # the other configs beside it (2^20-cell samples, best of 2), Gcells/s
    try:
        import tempfile
        W, tq = pkg.workloads, pkg.thermalqs
        ns = 1 << 20
        other = {}

        def best_of(fn, reps2=2):
            fn()
            b = 1e30
            for _ in range(reps2):
                t0 = time.perf_counter()
                fn()
                b = min(b, time.perf_counter() - t0)
            return b
        r1, t1 = W.ideal_inputs(ns)
        ig = ref.IdealGas(0.4, 7.16e6)
        other["ideal_P_then_e"] = ns / best_of(lambda: (
            ig.eval(L.P_FROM_RHO_T, r1, t1, nthreads=cores),
            ig.eval(L.E_FROM_RHO_T, r1, t1, nthreads=cores))) / 1e9
        rho_s, T_s = rho[:ns], T[:ns]
        sie_s = eos.eval(L.E_FROM_RHO_T, rho_s, T_s, nthreads=cores)
        lam_s = np.full((ns, 2), np.nan)

        def root_cold():
            lam_s.fill(np.nan)
            eos.eval(L.T_FROM_RHO_E, rho_s, sie_s, lam_s, nthreads=cores)
        other["spiner_rhot_T_rhoe_rootfind_cold"] = ns / best_of(root_cold) / 1e9
        rs = ref.SpinerRhoSie(raw)
        lam1 = np.full((ns, 1), np.nan)
        other["spiner_rhosie_T_rhoe"] = ns / best_of(
            lambda: rs.eval(L.T_FROM_RHO_E, rho_s, sie_s, lam1, nthreads=cores)) / 1e9
        with tempfile.TemporaryDirectory() as d:
            hp = pkg.tablegen.write_synthetic_helm_table(os.path.join(d, "helm.dat"))
            h = ref.Helmholtz(hp, L.HelmOptions(1, 1, 1, 1, 1, 1, 1))
            nh = 1 << 18
            hr, hT = W.helm_inputs(nh)
            lam3 = np.tile(np.array([4.0, 2.0, -1.0]), (nh, 1))
            he, hP, hcv, hb = (np.zeros(nh) for _ in range(4))
            mask = tq.specific_internal_energy | tq.pressure | tq.specific_heat | tq.bulk_modulus
            other["helmholtz_FillEos_rhoT"] = nh / best_of(
                lambda: h.fill(hr, hT, he, hP, hcv, hb, mask, lam3, nthreads=cores)) / 1e9
        out["other_configs_Gcells_per_s"] = other
    except Exception as exc:
        out["other_configs_error"] = repr(exc)
    # the reference's own get_sg_eos (Kokkos build only) on a 2^18-cell sample of the same inputs
    try:
        if ref.lib.ref_is_kokkos():
            r3 = ref.set_eos_3mat()
            nsg = 1 << 18
            a0, offs = sg_eos_inputs(
                pkg, nsg, lambda m, r_, t_: r3[m].eval(L.E_FROM_RHO_T, r_, t_, nthreads=cores))
            tsg = []
            for k in range(4):
                a1 = {kk: vv.copy(order="K") for kk, vv in a0.items()}
                t0 = time.perf_counter()
                ref.get_sg_eos(r3, [1, 2, 3], offs, **a1, input_int=1, mass_frac_cutoff=1e-8)
                if k >= 1:
                    tsg.append(time.perf_counter() - t0)
            out["get_sg_eos_rho_e_cells_per_s"] = nsg / float(np.median(tsg))
            out["get_sg_eos_sample"] = f"{nsg} cells, 1 warm-up + 3 timed calls, median, {cores} threads"
    except Exception as exc:
        out["get_sg_eos_error"] = repr(exc)
    return out


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the headline path (its
    Kokkos-OpenMP build) on all host threads, on the GPU arm's configuration: every step is the
    full 2^26-cell workload.  Rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import __graft_entry__ as g

    pkg = g.load_package()
    cores = host_threads()
    os.environ["OMP_NUM_THREADS"] = str(cores)
    ref, how = baseline_ref()
    if ref is None:
        print(json.dumps({"impl": "reference",
                          "unavailable": "oracle/_ref/libsgref*.so not built"}), flush=True)
        return
    L = pkg._lib
    raw = pkg.tablegen.vinet_sp5_tables()
    eos = ref.SpinerRhoT(raw)
    n = args.cells
    rho, T = pkg.workloads.spiner_inputs(n)
    lam = np.full((n, 2), np.nan)
    K, Wu = args.steps, max(args.warmup, 1)
    for _ in range(Wu):
        eos.eval(L.P_FROM_RHO_T, rho, T, lam, nthreads=cores)
    t0 = time.perf_counter()
    for _ in range(K):
        eos.eval(L.P_FROM_RHO_T, rho, T, lam, nthreads=cores)
    secs = time.perf_counter() - t0
    value = n * K / secs / 1e9
    world = args.gpus
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT,
        "n_gpus": args.gpus, "steps": K, "warmup": Wu, "ms_per_step": secs / K * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": "SpinerEOSDependsRhoT PressureFromDensityTemperature, 256x256 "
                               "NQT-log table from analytic Vinet, lambda (lRho,lT) write-back",
                   "cells_per_gpu": n, "global_cells": n * world,
                   "l2_policy": "inputs (1 GiB/GPU) larger than L2; no flush needed",
                   "parallelism": f"cells sharded contiguously over {world} GPU(s), tables replicated"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "reference",
                         "sample": f"{n} cells per step (one GPU's full shard of the workload) "
                                   f"on the host: {how}"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
