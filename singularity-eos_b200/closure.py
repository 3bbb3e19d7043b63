"""Batch PTE mixed-cell closure over the C ABI (sgx_pte_solve).

Mirrors closure/mixed_cell_models.hpp: MixParams (:43-78), PTESolverRhoT (:883-1193),
PTESolverPT (:1207-1446) driven by PTESolver (:2188-2291), launched over all cells at once
instead of the user's per-cell loop (test/test_pte.cpp:177-265).
"""
import ctypes as C

from . import _lib
from ._lib import check
from .eos import _ptr, _require_cuda_f64, _stream


def MixParams(**overrides):
    p = _lib.MixParams()
    _lib.load().sgx_mix_params_default(C.byref(p))
    for k, v in overrides.items():
        if not hasattr(p, k):
            raise AttributeError(k)
        setattr(p, k, v)
    return p


def test_pte_params():
    """The tolerances of test/test_pte.cpp:164-175."""
    return MixParams(pte_rel_tolerance_e=1e-16, pte_rel_tolerance_v=1e-16,
                     pte_rel_tolerance_p=1e-16, pte_abs_tolerance_p=1e-14,
                     pte_rel_tolerance_e_sufficient=1e-6, pte_rel_tolerance_v_sufficient=1e-6,
                     pte_rel_tolerance_p_sufficient=1e-6, pte_abs_tolerance_p_sufficient=1e-6,
                     iterate_t_guess=0, pte_small_step_thresh=1e-22, pte_small_step_tries=100)


def decode_status(words):
    """Unpack the 64-bit SolverStatus words (include/sgx_b200.h)."""
    w = words
    return {
        "converged": (w & 1) != 0,
        "slow_convergence_detected": ((w >> 1) & 1) != 0,
        "niter": (w >> 8) & 0xFFFF,
        "max_line_niter": (w >> 24) & 0xFF,
        "small_step_iters": (w >> 32) & 0xFFFF,
        "max_niter": (w >> 48) & 0xFFFF,
    }


def _handles(eos_list):
    arr = (C.c_void_p * len(eos_list))()
    for i, e in enumerate(eos_list):
        arr[i] = e.handle
    return arr


def pte_solve(solver, eos_list, rho, vfrac, sie, temp, press, sie_tot=None, vfrac_tot=None,
              Tguess=None, params=None, status=None, residual=None, stats=None, lambdas=None):
    """Solve all cells.  State tensors are [ncell][nmat] CUDA float64, updated in place.
    lambdas: [ncell][nmat][3] per-material caches (CacheAccessor layout,
    closure/mixed_cell_models.hpp:307-324) for mixtures with tabulated materials, or None."""
    _require_cuda_f64(rho, vfrac, sie, temp, press, sie_tot, vfrac_tot, Tguess, residual, lambdas)
    nmat = len(eos_list)
    ncell = rho.shape[0]
    assert rho.shape == (ncell, nmat)
    p = params if params is not None else MixParams()
    if lambdas is not None:
        assert lambdas.numel() == ncell * nmat * 3
    check(_lib.load().sgx_pte_solve_cached(
        int(solver), _handles(eos_list), nmat, ncell, _ptr(rho), _ptr(vfrac), _ptr(sie),
        _ptr(temp), _ptr(press), _ptr(sie_tot), _ptr(vfrac_tot), _ptr(Tguess), C.byref(p),
        _ptr(status), _ptr(residual), _ptr(stats), _ptr(lambdas), _stream()))


def PTESolverRhoT(eos_list, rho, vfrac, sie, temp, press, **kw):
    return pte_solve(_lib.PTE_RHOT, eos_list, rho, vfrac, sie, temp, press, **kw)


def PTESolverPT(eos_list, rho, vfrac, sie, temp, press, **kw):
    return pte_solve(_lib.PTE_PT, eos_list, rho, vfrac, sie, temp, press, **kw)


def PTESolverRhoU(eos_list, rho, vfrac, sie, temp, press, **kw):
    """PTESolverRhoU (closure/mixed_cell_models.hpp:1934-2183)."""
    return pte_solve(_lib.PTE_RHOU, eos_list, rho, vfrac, sie, temp, press, **kw)


def PTESolverFixedT(eos_list, T_true, rho, vfrac, sie, temp, press, **kw):
    """PTESolverFixedT (closure/mixed_cell_models.hpp:1462-1672): the cell temperatures
    `T_true` [ncell] are given; unknowns are the volume fractions."""
    return pte_solve(_lib.PTE_FIXED_T, eos_list, rho, vfrac, sie, temp, press, Tguess=T_true,
                     **kw)


def PTESolverFixedP(eos_list, P_true, rho, vfrac, sie, temp, press, **kw):
    """PTESolverFixedP (closure/mixed_cell_models.hpp:1674-1932): the cell pressures `P_true`
    [ncell] are given; unknowns are the volume fractions and the temperature.  params must have
    iterate_t_guess = 0 (see include/sgx_b200.h)."""
    return pte_solve(_lib.PTE_FIXED_P, eos_list, rho, vfrac, sie, temp, press, Tguess=P_true,
                     **kw)


def pte_solve_host(solver, eos_list, rho, vfrac, sie, temp, press, params=None, status=None,
                   residual=None, sie_tot=None, vfrac_tot=None, Tguess=None):
    """numpy arrays in host memory; staging happens inside the library."""
    def p(a):
        return None if a is None else a.ctypes.data

    nmat = len(eos_list)
    ncell = rho.shape[0]
    prm = params if params is not None else MixParams()
    check(_lib.load().sgx_pte_solve_host(
        int(solver), _handles(eos_list), nmat, ncell, p(rho), p(vfrac), p(sie), p(temp),
        p(press), p(sie_tot), p(vfrac_tot), p(Tguess), C.byref(prm), p(status), p(residual)))


def set_eos_3mat_analytic():
    """The three materials of test/pte_test_3mat_analytic.hpp:38-50."""
    from .eos import DavisProducts, DavisReactants, Gruneisen

    gr = Gruneisen(394000.0, 1.489, 0.0, 0.0, 2.02, 0.47, 8.93, 297.0, 1.0e6, 0.383e7)
    dr = DavisReactants(1.890, 4.115e10, 1.0e6, 297.0, 1.8e5, 4.6, 0.34, 0.56, 0.0, 0.4265,
                        0.001074e10)
    dp = DavisProducts(0.798311, 0.58, 1.35, 2.66182, 0.75419, 3.2e10, 0.001072e10)
    return [gr, dr, dp]


def get_sg_eos(eos_list, eos_offsets, offsets, press, pmax, vol, spvol, sie, temp, bmod, dpde, cv,
               frac_mass, frac_vol, frac_ie, frac_bmod=None, frac_dpde=None, frac_cv=None,
               input_int=1, mass_frac_cutoff=1e-16):
    """get_sg_eos (eos/get_sg_eos.cpp:36-58) on numpy HOST arrays, the reference's own memory
    contract: per-cell arrays [cell_dim], per-material arrays Fortran-ordered (cell_dim, nmat),
    1-based int32 offsets.  input_int: 0 / 1 (rho, e), -3 (rho, T), -1 (P, T)."""
    import numpy as np

    nmat = len(eos_offsets)
    ncell = len(offsets)
    cell_dim = press.shape[0]
    eo = np.ascontiguousarray(eos_offsets, dtype=np.int32)
    of = np.ascontiguousarray(offsets, dtype=np.int32)
    for a in (frac_mass, frac_vol, frac_ie, frac_bmod, frac_dpde, frac_cv):
        assert a is None or (a.flags.f_contiguous and a.shape == (cell_dim, nmat))

    def p(a):
        return None if a is None else a.ctypes.data

    check(_lib.load().sgx_get_sg_eos(
        nmat, ncell, cell_dim, int(input_int), eo.ctypes.data, _handles(eos_list), of.ctypes.data,
        p(press), p(pmax), p(vol), p(spvol), p(sie), p(temp), p(bmod), p(dpde), p(cv),
        p(frac_mass), p(frac_vol), p(frac_ie), p(frac_bmod), p(frac_dpde), p(frac_cv),
        float(mass_frac_cutoff)))
