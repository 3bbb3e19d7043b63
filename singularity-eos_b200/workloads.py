"""Synthetic inputs for the benchmark configurations of BASELINE.json.

Host (numpy) generators feed the parity tests -- the same bits go to the oracle and to the
CUDA library.  Device (torch) generators feed the timed runs, so 16 GB arrays never cross
PCIe.  Uniform variates come from rng.uniform_* (splitmix64 counter RNG; identical bits on
host and device).
"""
import numpy as np

from . import rng

EPS = 10 * np.finfo(np.float64).eps  # robust::EPS()

# seeds fixed by SURVEY 8d
SEED_IDEAL_RHO, SEED_IDEAL_T = 1, 2
SEED_SPINER_RHO, SEED_SPINER_T = 3, 4
SEED_ROOT_RHO, SEED_ROOT_T = 5, 6
SEED_PTE = 7
SEED_HELM_RHO, SEED_HELM_T = 8, 9

PTE_RHO = (8.93, 1.89, 2.5)  # test/pte_test_3mat_analytic.hpp:57-59
PTE_T = 600.0


def _u(xp, seed, start, n, device=None):
    if xp is np:
        return rng.uniform_np(seed, start, n)
    return rng.uniform_torch(seed, start, n, device)


def ideal_inputs(n, start=0, xp=np, device=None):
    """rho = 1e-2 + u (1e2 - 1e-2), T = 1e2 + u (1e4 - 1e2)  (test/profile_eos.cpp:52-55,81-82)"""
    rho = 1e-2 + _u(xp, SEED_IDEAL_RHO, start, n, device) * (1e2 - 1e-2)
    T = 1e2 + _u(xp, SEED_IDEAL_T, start, n, device) * (1e4 - 1e2)
    return rho, T


def _log_uniform(xp, u, lo, hi, inner=0.98):
    llo, lhi = np.log10(lo), np.log10(hi)
    pad = 0.5 * (1.0 - inner) * (lhi - llo)
    return 10.0 ** ((llo + pad) + u * ((lhi - pad) - (llo + pad)))


def spiner_inputs(n, start=0, xp=np, device=None, seeds=(SEED_SPINER_RHO, SEED_SPINER_T),
                  rho_range=(1.0, 30.0), T_range=(1.0, 1.0e5)):
    """rho, T log-uniform over the inner 98 % of the table."""
    rho = _log_uniform(xp, _u(xp, seeds[0], start, n, device), *rho_range)
    T = _log_uniform(xp, _u(xp, seeds[1], start, n, device), *T_range)
    return rho, T


def helm_inputs(n, start=0, xp=np, device=None):
    """log10 rho in U[-3, 9], log10 T in U[4, 10] (span of test/test_eos_helmholtz.cpp:68-69)"""
    lr = -3.0 + _u(xp, SEED_HELM_RHO, start, n, device) * 12.0
    lt = 4.0 + _u(xp, SEED_HELM_T, start, n, device) * 6.0
    return 10.0 ** lr, 10.0 ** lt


def pte_vfracs(n, start=0, xp=np, device=None, nmat=3):
    """Dirichlet(1,..,1) volume fractions as test/pte_test_3mat_analytic.hpp:61-79:
    v = -ln(1-u) + EPS, normalised, clamped to [EPS, 1].  Returns [n][nmat]."""
    u = _u(xp, SEED_PTE, start * nmat, n * nmat, device).reshape(n, nmat)
    if xp is np:
        v = -np.log1p(-u) + EPS
        v = v * (1.0 / v.sum(axis=1, keepdims=True))
        return np.ascontiguousarray(np.maximum(EPS, np.minimum(v, 1.0)))
    import torch

    v = -torch.log1p(-u) + EPS
    v = v * (1.0 / v.sum(dim=1, keepdim=True))
    return torch.clamp(v, min=EPS, max=1.0).contiguous()


def pte_state_np(n, sie_of_material, start=0):
    """Host PTE state: rho, vfrac, sie, temp, press [n][3].  `sie_of_material[m]` is
    eos[m].InternalEnergyFromDensityTemperature(rho_m, 600) supplied by the caller."""
    vfrac = pte_vfracs(n, start)
    rho = np.tile(np.array(PTE_RHO), (n, 1))
    sie = np.tile(np.asarray(sie_of_material, dtype=np.float64), (n, 1))
    temp = np.full((n, 3), PTE_T)
    press = np.zeros((n, 3))
    return rho, vfrac, sie, temp, press


def pte_state_torch(n, sie_of_material, device, start=0):
    import torch

    vfrac = pte_vfracs(n, start, xp=torch, device=device)
    rho = torch.tensor(PTE_RHO, dtype=torch.float64, device=device).repeat(n, 1).contiguous()
    sie = torch.as_tensor(np.asarray(sie_of_material), dtype=torch.float64,
                          device=device).repeat(n, 1).contiguous()
    temp = torch.full((n, 3), PTE_T, dtype=torch.float64, device=device)
    press = torch.zeros((n, 3), dtype=torch.float64, device=device)
    return rho, vfrac, sie, temp, press


def pte_test_state_mt19937(ntrial=100):
    """The exact 100 cells of test/test_pte.cpp:137-144: std::mt19937(123456789) +
    std::exponential_distribution<double>(1.0) as implemented by libstdc++
    (-log(1 - generate_canonical<double,53>(gen)) / lambda, two 32-bit draws per variate)."""
    class MT:
        def __init__(self, seed):
            self.mt = [0] * 624
            self.idx = 624
            self.mt[0] = seed & 0xFFFFFFFF
            for i in range(1, 624):
                self.mt[i] = (1812433253 * (self.mt[i - 1] ^ (self.mt[i - 1] >> 30)) + i) & 0xFFFFFFFF

        def next(self):
            if self.idx >= 624:
                for i in range(624):
                    y = (self.mt[i] & 0x80000000) | (self.mt[(i + 1) % 624] & 0x7FFFFFFF)
                    self.mt[i] = self.mt[(i + 397) % 624] ^ (y >> 1) ^ (0x9908B0DF if y & 1 else 0)
                self.idx = 0
            y = self.mt[self.idx]
            self.idx += 1
            y ^= y >> 11
            y ^= (y << 7) & 0x9D2C5680
            y ^= (y << 15) & 0xEFC60000
            y ^= y >> 18
            return y & 0xFFFFFFFF

    gen = MT(123456789)

    def canonical():
        # generate_canonical<double, 53>: k = 2 draws, sum += (g - min) * R^i, / R^k
        lo = float(gen.next())
        hi = float(gen.next())
        r = (lo + hi * 4294967296.0) / (4294967296.0 * 4294967296.0)
        if r >= 1.0:
            r = np.nextafter(1.0, 0.0)
        return r

    vf = np.empty((ntrial, 3))
    for n in range(ntrial):
        v = np.array([-np.log(1.0 - canonical()) / 1.0 + EPS for _ in range(3)])
        vsum = 0.0
        for x in v:
            vsum += x
        v = v * (1.0 / vsum)
        vf[n] = np.maximum(EPS, np.minimum(v, 1.0))
    return vf
