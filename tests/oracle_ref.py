"""TEST INFRASTRUCTURE: ctypes binding of oracle/_ref/libsgref.so, the unmodified reference
compiled for the host (oracle/ref_harness.cpp).  Only tests/, smoke() and bench.py's CPU
baseline legs may import this."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_LIB = os.path.join(ROOT, "oracle", "_ref", "libsgref.so")

_VP, _I64, _D = C.c_void_p, C.c_int64, C.c_double
KOKKOS_LIB = os.path.join(ROOT, "oracle", "_ref", "libsgref_kokkos.so")
_loaded = {}


def _pkg():
    import __graft_entry__ as g

    return g.load_package()


def load(path=None):
    """The parity oracle (default): the reference headers with PORTABILITY_STRATEGY_NONE, -O2."""
    path = path or REF_LIB
    if path in _loaded:
        return _loaded[path]
    if not os.path.exists(path):
        return None
    L = _pkg()._lib
    lib = C.CDLL(path)
    protos = {
        "ref_abi_version": (C.c_int, []),
        "ref_num_threads": (C.c_int, []),
        "ref_is_kokkos": (C.c_int, []),
        "ref_fastmath_eval": (None, [C.c_int, _VP, _VP, _I64]),
        "ref_regula_falsi_table": (C.c_int, [_VP, C.c_int, _D, _D, _D, _D, _D, _D, _D, _D,
                                             C.POINTER(_D)]),
        "ref_solve_Ax_b": (C.c_int, [C.c_int, _VP, _VP]),
        "ref_solve_Ax_b_batch": (None, [C.c_int, _VP, _VP, _VP, C.c_long]),
        "ref_sum_neumaier": (_D, [_VP, C.c_int]),
        "ref_eos_ideal": (_VP, [_D, _D]),
        "ref_eos_gruneisen": (_VP, [_D] * 11),
        "ref_eos_davis_reactants": (_VP, [_D] * 11),
        "ref_eos_davis_products": (_VP, [_D] * 7),
        "ref_eos_vinet": (_VP, [_D] * 8 + [_VP]),
        "ref_eos_spiner_rhot": (_VP, [C.POINTER(L.Sp5Tables), C.c_int, C.c_int]),
        "ref_eos_spiner_rhosie": (_VP, [C.POINTER(L.Sp5Tables), C.c_int, C.c_int]),
        "ref_eos_helmholtz": (_VP, [C.c_char_p, C.POINTER(L.HelmOptions)]),
        "ref_eos_free": (None, [_VP]),
        "ref_eos_modify": (_VP, [_VP, C.c_int, _VP]),
        "ref_eos_free_shallow": (None, [_VP]),
        "ref_eos_nlambda": (C.c_int, [_VP]),
        "ref_eos_preferred_input": (C.c_ulong, [_VP]),
        "ref_eos_maximum_density": (_D, [_VP]),
        "ref_eos_get_table": (_I64, [_VP, C.c_char_p, _VP, _I64]),
        "ref_eos_get_scalar": (C.c_int, [_VP, C.c_char_p, C.POINTER(_D)]),
        "ref_eos_eval": (C.c_int, [_VP, C.c_int, _VP, _VP, _VP, _I64, _VP, C.c_int, C.c_int]),
        "ref_eos_fill": (C.c_int, [_VP] * 7 + [_I64, C.c_ulong, _VP, C.c_int, C.c_int, C.c_int]),
        "ref_eos_eval_extra": (C.c_int, [_VP, C.c_int, _VP, _VP, _VP, _VP, _I64, _VP, C.c_int]),
        "ref_mix_params_default": (None, [C.POINTER(L.MixParams)]),
        "ref_pte_solve": (C.c_int, [C.c_int, C.POINTER(_VP), C.c_int, _I64] + [_VP] * 8 +
                          [C.POINTER(L.MixParams), _VP, _VP, _VP, C.c_int]),
        "ref_pte_solve_cached": (C.c_int, [C.c_int, C.POINTER(_VP), C.c_int, _I64] + [_VP] * 8 +
                                 [C.POINTER(L.MixParams), _VP, _VP, _VP, _VP, C.c_int]),
        "ref_get_sg_eos": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, _VP, C.POINTER(_VP),
                                     C.c_int, _VP] + [_VP] * 15 + [_D]),
        "ref_pte_fixedp_init_probe": (C.c_int, [C.POINTER(_VP), C.c_int] + [_VP] * 5 +
                                      [_D, C.POINTER(L.MixParams), _D]),
        "ref_pte_approx_T": (None, [C.POINTER(_VP), C.c_int, _I64, _VP, _VP, _VP, _VP]),
        "ref_bounds_log_grid": (C.c_int, [_D, _D, C.c_int, C.POINTER(L.Grid1D), C.POINTER(_D),
                                          _VP]),
    }
    for name, (res, args) in protos.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    _loaded[path] = Ref(lib)
    return _loaded[path]


def load_kokkos():
    """The TIMED baseline: the same harness built -O3 against the reference's vendored Kokkos
    (OpenMP backend, PORTABILITY_STRATEGY_KOKKOS): whole arrays go through one reference vector
    call.  Thread placement as BASELINE.md 3 (must be set before the OpenMP runtime starts)."""
    os.environ.setdefault("OMP_PROC_BIND", "spread")
    os.environ.setdefault("OMP_PLACES", "threads")
    ref = load(KOKKOS_LIB)
    if ref is not None:
        assert ref.lib.ref_is_kokkos() == 1
    return ref


def _p(a):
    return None if a is None else a.ctypes.data


class RefEos:
    def __init__(self, lib, h):
        if not h:
            raise RuntimeError("reference EOS construction failed")
        self.lib, self.h = lib, h

    def __del__(self):
        try:
            self.lib.ref_eos_free(self.h)
        except Exception:
            pass

    def modify(self, which, *params):
        """Variant::Modify: 0 ShiftedEOS(shift), 1 ScaledEOS(scale), 2 BilinearRampEOS(r0,a,b,c).
        The result shares the base's tables and keeps it alive."""
        p = np.zeros(4)
        p[:len(params)] = params
        h = self.lib.ref_eos_modify(self.h, int(which), _p(p))
        if not h:
            raise RuntimeError("modified type not in the oracle's variant")
        m = RefEosShallow(self.lib, h)
        m.base = self
        return m

    def eval(self, method, a, b, lambdas=None, nthreads=1):
        a = np.ascontiguousarray(a, dtype=np.float64)
        b = np.ascontiguousarray(b, dtype=np.float64)
        out = np.empty_like(a)
        stride = 0 if lambdas is None else lambdas.shape[-1]
        rc = self.lib.ref_eos_eval(self.h, method, _p(a), _p(b), _p(out), a.size, _p(lambdas),
                                   stride, nthreads)
        assert rc == 0
        return out

    def eval_extra(self, op, in1, in2=None, rho_guess=None, lambdas=None):
        """ref_eos_eval_extra: returns out1, or (rho, sie) for op 6."""
        in1 = np.ascontiguousarray(in1, dtype=np.float64)
        in2 = None if in2 is None else np.ascontiguousarray(in2, dtype=np.float64)
        out1 = np.zeros_like(in1) if rho_guess is None else np.array(rho_guess, dtype=np.float64)
        out2 = np.zeros_like(in1)
        stride = 0 if lambdas is None else lambdas.shape[-1]
        rc = self.lib.ref_eos_eval_extra(self.h, op, _p(in1), _p(in2), _p(out1), _p(out2),
                                         in1.size, _p(lambdas), stride)
        if rc != 0:
            raise RuntimeError("the reference threw")
        return (out1, out2) if op == 6 else out1

    def fill(self, rho, T, e, P, cv, bmod, output, lambdas=None, az_only=False, nthreads=1):
        n = rho.size
        stride = 0 if lambdas is None else lambdas.shape[-1]
        rc = self.lib.ref_eos_fill(self.h, _p(rho), _p(T), _p(e), _p(P), _p(cv), _p(bmod), n,
                                   int(output), _p(lambdas), stride, int(az_only), nthreads)
        assert rc == 0

    def get_table(self, name):
        n = self.lib.ref_eos_get_table(self.h, name.encode(), None, 0)
        if n < 0:
            raise KeyError(name)
        out = np.empty(n)
        self.lib.ref_eos_get_table(self.h, name.encode(), _p(out), n)
        return out

    def get_scalar(self, name):
        v = _D()
        if self.lib.ref_eos_get_scalar(self.h, name.encode(), C.byref(v)) != 0:
            raise KeyError(name)
        return v.value

    def nlambda(self):
        return self.lib.ref_eos_nlambda(self.h)

    def MaximumDensity(self):
        return self.lib.ref_eos_maximum_density(self.h)



class RefEosShallow(RefEos):
    """A modified copy: deleting it must not Finalize() the tables it shares with its base."""

    def __del__(self):
        try:
            self.lib.ref_eos_free_shallow(self.h)
        except Exception:
            pass

class Ref:
    def __init__(self, lib):
        self.lib = lib

    def num_threads(self):
        return self.lib.ref_num_threads()

    def fastmath(self, which, x):
        x = np.ascontiguousarray(x, dtype=np.float64)
        y = np.empty_like(x)
        self.lib.ref_fastmath_eval(which, _p(x), _p(y), x.size)
        return y

    def regula_falsi_table(self, y, xmin, xmax, ytarget, guess, a, b, xtol, ytol):
        y = np.ascontiguousarray(y, dtype=np.float64)
        root = _D()
        st = self.lib.ref_regula_falsi_table(_p(y), y.size, xmin, xmax, ytarget, guess, a, b,
                                             xtol, ytol, C.byref(root))
        return st, root.value

    def solve_Ax_b(self, a, b):
        a = np.array(a, dtype=np.float64).copy()
        b = np.array(b, dtype=np.float64).copy()
        ok = self.lib.ref_solve_Ax_b(b.size, _p(a), _p(b))
        return ok, b

    def solve_Ax_b_batch(self, n, a, b):
        """a [batch][n*n], b [batch][n] (copies); returns (ok[batch], x[batch][n])."""
        a = np.array(a, dtype=np.float64).copy()
        b = np.array(b, dtype=np.float64).copy()
        ok = np.zeros(b.shape[0], dtype=np.int32)
        self.lib.ref_solve_Ax_b_batch(n, _p(a), _p(b), ok.ctypes.data_as(_VP), b.shape[0])
        return ok, b

    def sum_neumaier(self, x):
        x = np.ascontiguousarray(x, dtype=np.float64)
        return self.lib.ref_sum_neumaier(_p(x), x.size)

    def IdealGas(self, gm1, Cv):
        return RefEos(self.lib, self.lib.ref_eos_ideal(gm1, Cv))

    def Gruneisen(self, *a, rho_max=-1.0):
        return RefEos(self.lib, self.lib.ref_eos_gruneisen(*a, rho_max))

    def DavisReactants(self, *a):
        return RefEos(self.lib, self.lib.ref_eos_davis_reactants(*a))

    def DavisProducts(self, *a):
        return RefEos(self.lib, self.lib.ref_eos_davis_products(*a))

    def Vinet(self, rho0, T0, B0, BP0, A0, Cv0, E0, S0, d2to40=None):
        d = np.zeros(39) if d2to40 is None else np.ascontiguousarray(d2to40, dtype=np.float64)
        return RefEos(self.lib, self.lib.ref_eos_vinet(rho0, T0, B0, BP0, A0, Cv0, E0, S0, _p(d)))

    def SpinerRhoT(self, raw, reproducible=False, pmin_vapor_dome=False):
        return RefEos(self.lib, self.lib.ref_eos_spiner_rhot(raw.byref(), int(reproducible),
                                                             int(pmin_vapor_dome)))

    def SpinerRhoSie(self, raw, reproducible=False, pmin_vapor_dome=False):
        return RefEos(self.lib, self.lib.ref_eos_spiner_rhosie(raw.byref(), int(reproducible),
                                                               int(pmin_vapor_dome)))

    def Helmholtz(self, path, opt):
        return RefEos(self.lib, self.lib.ref_eos_helmholtz(path.encode(), C.byref(opt)))

    def set_eos_3mat(self):
        return [self.Gruneisen(394000.0, 1.489, 0.0, 0.0, 2.02, 0.47, 8.93, 297.0, 1.0e6, 0.383e7),
                self.DavisReactants(1.890, 4.115e10, 1.0e6, 297.0, 1.8e5, 4.6, 0.34, 0.56, 0.0,
                                    0.4265, 0.001074e10),
                self.DavisProducts(0.798311, 0.58, 1.35, 2.66182, 0.75419, 3.2e10, 0.001072e10)]

    def mix_params_default(self):
        p = _pkg()._lib.MixParams()
        self.lib.ref_mix_params_default(C.byref(p))
        return p

    def pte_solve(self, solver, eos_list, rho, vfrac, sie, temp, press, params=None,
                  sie_tot=None, vfrac_tot=None, Tguess=None, nthreads=1, lambdas=None):
        """In-place on the numpy state arrays; returns (status words, residual, Tguess).
        lambdas: [ncell][nmat][3] CacheAccessor storage (in/out) or None (NullIndexer)."""
        nmat = len(eos_list)
        ncell = rho.shape[0]
        hs = (_VP * nmat)(*[e.h for e in eos_list])
        status = np.zeros(ncell, dtype=np.uint64)
        resid = np.zeros(ncell)
        tg = np.zeros(ncell)
        rc = self.lib.ref_pte_solve_cached(
            int(solver), hs, nmat, ncell, _p(rho), _p(vfrac), _p(sie), _p(temp), _p(press),
            _p(sie_tot), _p(vfrac_tot), _p(Tguess), None if params is None else C.byref(params),
            _p(status), _p(resid), _p(tg), _p(lambdas), nthreads)
        assert rc == 0
        return status, resid, tg

    def pte_fixedp_init_probe(self, eos_list, rho, vfrac, sie, temp, press, P, params, fill):
        """PTESolverFixedP::Init() of ONE cell, the solver object constructed over memory that
        holds the double `fill`; the state arrays ([nmat]) are updated in place."""
        hs = (_VP * len(eos_list))(*[e.h for e in eos_list])
        return self.lib.ref_pte_fixedp_init_probe(hs, len(eos_list), _p(rho), _p(vfrac), _p(sie),
                                                  _p(temp), _p(press), float(P),
                                                  C.byref(params), float(fill))

    def get_sg_eos(self, eos_list, eos_offsets, offsets, press, pmax, vol, spvol, sie, temp,
                   bmod, dpde, cv, frac_mass, frac_vol, frac_ie, frac_bmod=None, frac_dpde=None,
                   frac_cv=None, input_int=1, mass_frac_cutoff=1e-16):
        """The reference's own get_sg_eos (Kokkos build of the harness only); numpy host arrays,
        per-material arrays Fortran-ordered (cell_dim, nmat), 1-based int32 offsets."""
        nmat, ncell, cell_dim = len(eos_offsets), len(offsets), press.shape[0]
        eo = np.ascontiguousarray(eos_offsets, dtype=np.int32)
        of = np.ascontiguousarray(offsets, dtype=np.int32)
        hs = (_VP * len(eos_list))(*[e.h for e in eos_list])
        rc = self.lib.ref_get_sg_eos(
            nmat, ncell, cell_dim, int(input_int), _p(eo), hs, len(eos_list), _p(of), _p(press),
            _p(pmax), _p(vol), _p(spvol), _p(sie), _p(temp), _p(bmod), _p(dpde), _p(cv),
            _p(frac_mass), _p(frac_vol), _p(frac_ie), _p(frac_bmod), _p(frac_dpde), _p(frac_cv),
            float(mass_frac_cutoff))
        assert rc == 0, rc

    def bounds_log_grid(self, xmin, xmax, n):
        g = _pkg()._lib.Grid1D()
        off = _D()
        nodes = np.empty(n)
        self.lib.ref_bounds_log_grid(xmin, xmax, n, C.byref(g), C.byref(off), _p(nodes))
        return g, off.value, nodes
