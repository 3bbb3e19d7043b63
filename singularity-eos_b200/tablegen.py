"""Synthetic table generators for the benchmark configurations.

The reference has no path that tabulates an analytic EOS into a Spiner EOS (its .sp5 files
come from sesame2spiner + proprietary EOSPAC), and the shipped Helmholtz table is absent from
the checkout.  These generators produce inputs of the same *shape and format*:

* `vinet_sp5_tables`  -- the raw contents of one .sp5 material (every dataset
  SpinerEOSDependsRhoT / RhoSie::loadDataboxes_ reads, eos_spiner_rho_temp.hpp:545-583,
  eos_spiner_rho_sie.hpp:434-488) tabulated from the analytic Vinet EOS
  (eos/eos_vinet.hpp:228-292) on NQT-log grids built like table_utils::Bounds
  (base/spiner_table_utils.hpp:56-70,190-206).
* `write_synthetic_helm_table` -- a 4-block ASCII file in the layout
  Helmholtz reads (eos_helmholtz.hpp:962-975).

This is data generation, not the hot path: the same raw tables are handed to the CUDA library
and to the oracle, each of which then performs its own load-time fix-ups.
"""
import ctypes as C

import numpy as np

from . import _lib

# ---------------------------------------------------------------------------------------
# NQT order-2 logs on numpy int64 (base/fast-math/logs.hpp:188-214) -- used for grid bounds
# ---------------------------------------------------------------------------------------
_ONE_AS_INT = (1 << 62) - (1 << 52)
_MANT = (1 << 52) - 1
_LOW = (1 << 26) - 1
_SCALE_DOWN = 2.22044604925031e-16
_SCALE_UP = float(1 << 52)


def fast_lg(x):
    x = np.asarray(x, dtype=np.float64)
    xi = x.view(np.int64) - np.int64(_ONE_AS_INT)
    frac = xi & np.int64(_MANT)
    fh = frac >> np.int64(26)
    fl = frac & np.int64(_LOW)
    fsq = fh * fh + ((fh * fl) >> np.int64(25))
    num = frac - fsq
    q = np.where(num >= 0, num // 3, -((-num) // 3))  # C++ truncating division
    return (xi + q).astype(np.float64) * _SCALE_DOWN


def fast_pow2(x):
    x = np.asarray(x, dtype=np.float64)
    a, b, c = np.int64(9007199254740992), 67108864.0, np.int64(18014398509481984)
    xi = np.trunc(x * _SCALE_UP).astype(np.int64)
    frac = xi & np.int64(_MANT)
    fsqrt = np.trunc(b * np.sqrt((c - np.int64(3) * frac).astype(np.float64))).astype(np.int64)
    return (xi + a - fsqrt - frac + np.int64(_ONE_AS_INT)).view(np.float64)


def fast_log10(x):
    return 0.301029995663981195 * fast_lg(x)


def fast_pow10(x):
    return fast_pow2(3.321928094887362626 * np.asarray(x, dtype=np.float64))


def log_bounds(xmin, xmax, n):
    """Single-patch Bounds(min, max, N, convertToLog=True): returns (Grid1D, offset, nodes)."""
    eps = np.finfo(np.float64).eps
    offset = 0.0
    if xmin <= 0:
        offset = 1.1 * abs(xmin) + 10 * eps
    lmin = float(fast_log10(np.abs(np.float64(xmin + offset))))
    lmax = float(fast_log10(np.abs(np.float64(xmax + offset))))
    dx = (lmax - lmin) / float(n - 1)
    g = _lib.Grid1D()
    g.ngrids = 1
    g.npoints[0] = n
    g.min[0], g.max[0], g.dx[0] = lmin, lmax, dx
    i = np.arange(n, dtype=np.float64)
    lnodes = i * dx + lmin  # RegularGrid1D::x
    nodes = np.minimum(xmax, np.maximum(xmin, fast_pow10(lnodes) - offset))  # Bounds::log2lin
    return g, offset, nodes


def piecewise_log_bounds(xmin, xmax, npoints=(48, 160, 48), breaks=(0.3, 0.7)):
    """Multi-patch axis = Spiner::PiecewiseGrid1D (utils/spiner/spiner/piecewise_grid_1d.hpp:
    47-68): RegularGrid1D patches that meet at exactly one point, so the junction node is stored
    twice and every table carries sum(npoints) entries per axis.  This is the coarse-fine-coarse
    shape sesame2spiner writes (base/spiner_table_utils.hpp:72-135).  `breaks` are the patch
    junctions as fractions of the log range.  Returns (Grid1D, offset, nodes)."""
    eps = np.finfo(np.float64).eps
    offset = 0.0
    if xmin <= 0:
        offset = 1.1 * abs(xmin) + 10 * eps
    lmin = float(fast_log10(np.abs(np.float64(xmin + offset))))
    lmax = float(fast_log10(np.abs(np.float64(xmax + offset))))
    edges = [lmin] + [lmin + f * (lmax - lmin) for f in breaks] + [lmax]
    assert len(npoints) == len(edges) - 1 <= 3
    g = _lib.Grid1D()
    g.ngrids = len(npoints)
    lnodes = []
    for k, n in enumerate(npoints):
        lo, hi = edges[k], edges[k + 1]
        dx = (hi - lo) / float(n - 1)
        g.npoints[k] = n
        g.min[k], g.max[k], g.dx[k] = lo, hi, dx
        lnodes.append(np.arange(n, dtype=np.float64) * dx + lo)  # RegularGrid1D::x
    lnodes = np.concatenate(lnodes)
    nodes = np.minimum(xmax, np.maximum(xmin, fast_pow10(lnodes) - offset))
    return g, offset, nodes


# ---------------------------------------------------------------------------------------
# Vinet (eos/eos_vinet.hpp:204-292)
# ---------------------------------------------------------------------------------------
class Vinet:
    def __init__(self, rho0=8.93, T0=298.0, B0=1.3448466e12, BP0=4.956, A0=5.19245e-05,
                 Cv0=0.383e7, E0=0.0, S0=5.05e8, d2to40=None):
        self.rho0, self.T0, self.B0, self.BP0 = rho0, T0, B0, BP0
        self.A0, self.Cv0, self.E0, self.S0 = A0, Cv0, E0, S0
        d = np.zeros(39) if d2to40 is None else np.asarray(d2to40, dtype=np.float64)
        self.d = d
        vip = np.zeros(43)
        vip[0] = B0 / rho0 + A0 * A0 * B0 * B0 * T0 / (rho0 * rho0 * Cv0)
        vip[1] = 1.5 * (BP0 - 1.0)
        vip[2], vip[3] = 1.0, 0.0
        vip[4:] = d
        for ind in range(41, 1, -1):
            vip[ind] = vip[ind] - ind / vip[1] * vip[ind + 1]
        self.vip = vip

    def eval(self, rho, T):
        """Returns e, P, dPdrho_T, dPdT_rho, dedT_rho, dedrho_T (Vinet_F_DT_func)."""
        rho = np.asarray(rho, dtype=np.float64)
        T = np.asarray(T, dtype=np.float64)
        vip, d = self.vip, self.d
        x = np.cbrt(self.rho0 / rho)
        x2inv = 1.0 / (x * x)
        onemx = 1.0 - x
        eta1mx = vip[1] * onemx
        ex = np.exp(eta1mx)
        sumP = np.zeros_like(x)
        sumB = np.zeros_like(x)
        sumE = np.zeros_like(x)
        for ind in range(40, 1, -1):
            sumP = d[ind - 2] + onemx * sumP
            sumB = d[ind - 2] * ind + onemx * sumB
            sumE = vip[2 + ind] + onemx * sumE
        for ind in (1, 0):
            sumE = vip[2 + ind] + onemx * sumE
        sumP = 1.0 + sumP * (onemx * onemx)
        sumB = sumB * onemx
        sumE = vip[1] * onemx * sumE
        AB = self.A0 * self.B0
        energy = (9.0 * self.B0 / (vip[1] * vip[1] * self.rho0) * (vip[2] - ex * (vip[2] - sumE))
                  - AB * (self.T0 / self.rho0 - self.T0 / rho))
        pressure = 3.0 * self.B0 * x2inv * onemx * ex * sumP
        tmp = (1.0 + onemx * (vip[1] * x + 1.0)) * sumP + x * onemx * sumB
        tmp = self.B0 / self.rho0 * x * ex * tmp
        energy = energy + self.Cv0 * (T - self.T0) + self.E0
        pressure = pressure + AB * (T - self.T0)
        dpdrho = tmp + 0.0 * T
        dpdt = AB + 0.0 * pressure
        dedt = self.Cv0 + 0.0 * pressure
        dedrho = (pressure - T * AB) / (rho * rho)
        return energy, pressure, dpdrho, dpdt, dedt, dedrho

    def T_from_rho_e(self, rho, e):
        e0, *_ = self.eval(rho, np.full_like(np.asarray(rho, dtype=np.float64), self.T0))
        return self.T0 + (e - e0 - 0.0) / self.Cv0


class Sp5Raw:
    """Owns the numpy arrays a sgx_sp5_tables struct points into."""

    def __init__(self):
        self.struct = _lib.Sp5Tables()
        self.arrays = {}

    def set(self, name, arr):
        a = np.ascontiguousarray(arr, dtype=np.float64)
        self.arrays[name] = a
        setattr(self.struct, name, a.ctypes.data_as(C.POINTER(C.c_double)))

    def byref(self):
        return C.byref(self.struct)


def _derived(rho, e, P, dpdrho_T, dpdt, cv, dedrho_T):
    out = {}
    out["dPdRho"] = dpdrho_T - dpdt * dedrho_T / cv  # (dP/drho)_e
    out["dPdE"] = dpdt / cv  # (dP/de)_rho
    out["dTdRho"] = -dedrho_T / cv  # (dT/drho)_e
    out["dTdE"] = 1.0 / cv
    out["dEdRho"] = dedrho_T
    out["dEdT"] = cv
    out["bMod"] = rho * out["dPdRho"] + (P / rho) * out["dPdE"]
    return out


def vinet_sp5_tables(nrho=256, nT=256, nE=256, rho_min=1.0, rho_max=30.0, T_min=1.0,
                     T_max=1.0e5, vinet=None, piecewise=False):
    """Vinet copper on 256x256 single-patch NQT-log grids, or
    (`piecewise=True`) on 3-patch PiecewiseGrid1D axes with the same node totals."""
    v = vinet or Vinet()
    raw = Sp5Raw()
    if piecewise:
        split = lambda n: (n // 4 - 16, n - 2 * (n // 4 - 16), n // 4 - 16)
        gR, offR, rho_n = piecewise_log_bounds(rho_min, rho_max, split(nrho), (0.25, 0.6))
        gT, offT, T_n = piecewise_log_bounds(T_min, T_max, split(nT), (0.3, 0.7))
    else:
        gR, offR, rho_n = log_bounds(rho_min, rho_max, nrho)
        gT, offT, T_n = log_bounds(T_min, T_max, nT)
    R, TT = np.meshgrid(rho_n, T_n, indexing="ij")  # [nrho][nT], T fastest
    e, P, dpdr, dpdt, cv, dedr = v.eval(R, TT)
    der = _derived(R, e, P, dpdr, dpdt, cv, dedr)
    raw.set("P", P)
    raw.set("sie", e)
    for k in ("bMod", "dPdRho", "dPdE", "dTdRho", "dTdE", "dEdRho", "dEdT"):
        raw.set(k, der[k])
    # cold curve = T=0 isotherm (sesame2spiner/io_eospac.cpp:250)
    e0, P0, dpdr0, dpdt0, cv0, dedr0 = v.eval(rho_n, np.zeros_like(rho_n))
    der0 = _derived(rho_n, e0, P0, dpdr0, dpdt0, cv0, dedr0)
    raw.set("PCold", P0)
    raw.set("sieCold", e0)
    raw.set("bModCold", der0["bMod"])
    raw.set("dPdRhoCold", der0["dPdRho"])
    # (rho, sie) set
    if piecewise:
        gE, offE, e_n = piecewise_log_bounds(float(e.min()), float(e.max()), split(nE),
                                             (0.35, 0.65))
    else:
        gE, offE, e_n = log_bounds(float(e.min()), float(e.max()), nE)
    R2, EE = np.meshgrid(rho_n, e_n, indexing="ij")
    T2 = v.T_from_rho_e(R2, EE)
    e2, P2, dpdr2, dpdt2, cv2, dedr2 = v.eval(R2, T2)
    der2 = _derived(R2, e2, P2, dpdr2, dpdt2, cv2, dedr2)
    raw.set("T_re", T2)
    raw.set("P_re", P2)
    for k in ("bMod", "dPdRho", "dPdE", "dTdRho", "dTdE", "dEdRho"):
        raw.set(k + "_re", der2[k])
    s = raw.struct
    s.lrho, s.lT, s.lE = gR, gT, gE
    s.rho_offset, s.T_offset, s.sie_offset = offR, offT, offE
    s.normal_density = v.rho0
    s.mean_atomic_mass, s.mean_atomic_number = 63.546, 29.0
    raw.nodes = {"rho": rho_n, "T": T_n, "e": e_n}
    return raw


# ---------------------------------------------------------------------------------------
# Synthetic Helmholtz free-energy table (SURVEY 0 finding 4: data/helmholtz/helm_table.dat is
# not in the checkout).  Same 4-block ASCII layout the reference reads
# (eos_helmholtz.hpp:962-975): 101 (T) x 271 (rho) rows per block, T outer / rho inner;
# block 1 = f fd ft fdd ftt fdt fddt fdtt fddtt, block 2 = dpdf dpdfd dpdft dpdfdt,
# block 3 = ef efd eft efdt, block 4 = xf xfd xft xfdt.
#
# Physics: a non-degenerate ideal electron gas plus a cold (T-independent) Fermi-like term,
# with every mixed derivative exact, so (de/dT)_rho > 0 everywhere and the reference's Newton
# iteration on e(T) converges:
#   f(d,T) = A T (L - 1) + B d^(2/3),  L = ln d - 1.5 ln T + C,  A = N_A k_B,  d = rho*Ye
#   => p = d^2 f_d = A T d + (2/3) B d^(5/3),   e = f - T f_T = 1.5 A T + B d^(2/3)
# ---------------------------------------------------------------------------------------
HELM_NT, HELM_NRHO = 101, 271
HELM_LT = (3.0, 13.0)
HELM_LRHO = (-12.0, 15.0)
_HELM_KB, _HELM_NA = 1.3806504e-16, 6.02214199e23
_HELM_LN10 = 2.30258509299405e+00


def synthetic_helm_arrays(B=1.0e13, C=-10.0):
    """Returns the 21 node tables [NT][NRHO] in file-column order, grouped per block."""
    A = _HELM_KB * _HELM_NA
    dlr = (HELM_LRHO[1] - HELM_LRHO[0]) / float(HELM_NRHO - 1)
    dlt = (HELM_LT[1] - HELM_LT[0]) / float(HELM_NT - 1)
    # the node coordinates the reference itself uses: pow10(range.x(i)) = exp(ln10 * x)
    d = np.exp(_HELM_LN10 * (np.arange(HELM_NRHO) * dlr + HELM_LRHO[0]))[None, :]
    T = np.exp(_HELM_LN10 * (np.arange(HELM_NT) * dlt + HELM_LT[0]))[:, None]
    one = np.ones((HELM_NT, HELM_NRHO))
    L = np.log(d) - 1.5 * np.log(T) + C
    d23 = d ** (2.0 / 3.0)
    f = A * T * (L - 1.0) + B * d23
    fd = A * T / d + (2.0 / 3.0) * B * d23 / d
    ft = A * (L - 2.5) * one
    fdd = -A * T / d**2 - (2.0 / 9.0) * B * d23 / d**2
    ftt = -1.5 * A / T * one
    fdt = A / d * one
    fddt = -A / d**2 * one
    fdtt = 0.0 * one
    fddtt = 0.0 * one
    blk1 = [f, fd, ft, fdd, ftt, fdt, fddt, fdtt, fddtt]
    dpdf = A * T + (10.0 / 9.0) * B * d23
    dpdfd = (20.0 / 27.0) * B * d23 / d * one
    dpdft = A * one
    dpdfdt = 0.0 * one
    blk2 = [dpdf, dpdfd, dpdft, dpdfdt]
    blk3 = [L * one, 1.0 / d * one, -1.5 / T * one, 0.0 * one]
    blk4 = [_HELM_NA * d * one, _HELM_NA * one, 0.0 * one, 0.0 * one]
    return [blk1, blk2, blk3, blk4]


def write_synthetic_helm_table(path, B=1.0e13, C=-10.0):
    """Write the synthetic table as text with 17 significant digits (round-trips exactly)."""
    with open(path, "w") as fh:
        for blk in synthetic_helm_arrays(B, C):
            cols = np.stack([np.broadcast_to(a, (HELM_NT, HELM_NRHO)).reshape(-1) for a in blk],
                            axis=1)
            np.savetxt(fh, cols, fmt="%.17e")
    return path


def write_sp5_raw(raw, path):
    """Dump a Sp5Raw for a C++ consumer (tests/cpp/test_plugin.cpp): the sgx_sp5_tables struct
    bytes (pointer members meaningless), then for every table member in struct order an int64
    element count (0 = NULL) followed by the doubles."""
    with open(path, "wb") as fh:
        fh.write(b"SGXSP5R1")
        fh.write(bytes(raw.struct))
        for name in _lib.SP5_TABLE_NAMES:
            a = raw.arrays.get(name)
            n = 0 if a is None else a.size
            fh.write(np.int64(n).tobytes())
            if n:
                fh.write(np.ascontiguousarray(a, dtype=np.float64).tobytes())
    return path
