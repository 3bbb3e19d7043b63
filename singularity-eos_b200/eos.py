"""Host-side mirror of the reference's EOS vector interface for the hot path, over the C ABI.

Class and method names follow the reference (eos/eos_ideal.hpp, eos_gruneisen.hpp,
eos_davis.hpp, eos_spiner_rho_temp.hpp, eos_spiner_rho_sie.hpp, eos_helmholtz.hpp); the vector
methods take the same argument order as the reference's raw-pointer overloads
(eos/eos_base.hpp:196-203):  NAME(in1, in2, out, num, lambdas).
Arrays are torch CUDA float64 tensors (torch is the device-memory / stream plumbing only);
calls are asynchronous on torch's current stream, like the reference's Kokkos path.
"""
import ctypes as C

from . import _lib
from ._lib import check, thermalqs  # noqa: F401


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream():
    import torch

    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _require_cuda_f64(*tensors):
    import torch

    for t in tensors:
        if t is None:
            continue
        if not (isinstance(t, torch.Tensor) and t.is_cuda and t.dtype == torch.float64
                and t.is_contiguous()):
            raise TypeError("sgx_b200 expects contiguous CUDA float64 tensors "
                            "(there is no CPU path)")


class EosBase:
    """Common vector surface (SG_EOS_VEC_2IN_1OUT, eos/eos_base.hpp:169-230)."""

    def __init__(self, handle):
        self._h = handle
        self._lib = _lib.load()

    # -- lifetime: Finalize() -------------------------------------------------------------
    def Finalize(self):
        if self._h is not None:
            self._lib.sgx_eos_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.Finalize()
        except Exception:
            pass

    def GetOnDevice(self):  # tables are already device-resident
        return self

    @property
    def handle(self):
        return self._h

    def type(self):
        """enum sgx_eos_kind of the handle (SGX_EOS_MODIFIED for the modifier wrappers)."""
        return self._lib.sgx_eos_type(self._h)

    def nlambda(self):
        return self._lib.sgx_eos_nlambda(self._h)

    def PreferredInput(self):
        return self._lib.sgx_eos_preferred_input(self._h)

    def MaximumDensity(self):
        return self._lib.sgx_eos_maximum_density(self._h)

    # -- Serialize / DeSerialize (eos/eos_base.hpp:756-798) -------------------------------
    def SerializedSizeInBytes(self):
        return int(self._lib.sgx_eos_serialized_size(self._h))

    def Serialize(self):
        """The handle as a self-contained bytes object (parameters / raw tables / options)."""
        n = self.SerializedSizeInBytes()
        buf = C.create_string_buffer(n)
        check(self._lib.sgx_eos_serialize(self._h, buf, n))
        return buf.raw

    @classmethod
    def DeSerialize(cls, data):
        """An equivalent EOS on the current device, rebuilt from Serialize()'s bytes."""
        lib = _lib.load()
        h = C.c_void_p()
        check(lib.sgx_eos_deserialize(data, len(data), C.byref(h)))
        obj = EosBase.__new__(cls)
        EosBase.__init__(obj, h)
        return obj

    def scratch_size(self, method, nelements):
        return self._lib.sgx_eos_scratch_bytes(self._h, 0, nelements)

    def _lambda_stride(self, lambdas, n):
        """Row stride of a lambda array: shape[-1] for [n][stride] arrays; a flat array holds
        n rows of nlambda() values.  The array must cover n rows."""
        if lambdas is None:
            return 0
        ndim = lambdas.dim() if hasattr(lambdas, "dim") else lambdas.ndim
        stride = int(lambdas.shape[-1]) if ndim > 1 else max(1, self.nlambda())
        size = lambdas.numel() if hasattr(lambdas, "numel") else lambdas.size
        if size < n * stride:
            raise ValueError(f"lambda array holds {size} values, {n} rows of stride {stride} "
                             f"need {n * stride}")
        return stride

    def _eval(self, method, in1, in2, out, num, lambdas):
        _require_cuda_f64(in1, in2, out, lambdas)
        n = in1.numel() if num is None else int(num)
        stride = self._lambda_stride(lambdas, n)
        check(self._lib.sgx_eos_eval(self._h, method, _ptr(in1), _ptr(in2), _ptr(out), n,
                                     _ptr(lambdas), stride, _stream()))
        return out

    def PressureFromDensityTemperature(self, rhos, temperatures, pressures, num=None,
                                       lambdas=None):
        return self._eval(_lib.P_FROM_RHO_T, rhos, temperatures, pressures, num, lambdas)

    def InternalEnergyFromDensityTemperature(self, rhos, temperatures, sies, num=None,
                                             lambdas=None):
        return self._eval(_lib.E_FROM_RHO_T, rhos, temperatures, sies, num, lambdas)

    def TemperatureFromDensityInternalEnergy(self, rhos, sies, temperatures, num=None,
                                             lambdas=None):
        return self._eval(_lib.T_FROM_RHO_E, rhos, sies, temperatures, num, lambdas)

    def PressureFromDensityInternalEnergy(self, rhos, sies, pressures, num=None, lambdas=None):
        return self._eval(_lib.P_FROM_RHO_E, rhos, sies, pressures, num, lambdas)

    def SpecificHeatFromDensityTemperature(self, rhos, temperatures, cvs, num=None,
                                           lambdas=None):
        return self._eval(_lib.CV_FROM_RHO_T, rhos, temperatures, cvs, num, lambdas)

    def SpecificHeatFromDensityInternalEnergy(self, rhos, sies, cvs, num=None, lambdas=None):
        return self._eval(_lib.CV_FROM_RHO_E, rhos, sies, cvs, num, lambdas)

    def BulkModulusFromDensityTemperature(self, rhos, temperatures, bmods, num=None,
                                          lambdas=None):
        return self._eval(_lib.BMOD_FROM_RHO_T, rhos, temperatures, bmods, num, lambdas)

    def BulkModulusFromDensityInternalEnergy(self, rhos, sies, bmods, num=None, lambdas=None):
        return self._eval(_lib.BMOD_FROM_RHO_E, rhos, sies, bmods, num, lambdas)

    def GruneisenParamFromDensityTemperature(self, rhos, temperatures, gm1s, num=None,
                                             lambdas=None):
        return self._eval(_lib.GAMMA_FROM_RHO_T, rhos, temperatures, gm1s, num, lambdas)

    def GruneisenParamFromDensityInternalEnergy(self, rhos, sies, gm1s, num=None,
                                                lambdas=None):
        return self._eval(_lib.GAMMA_FROM_RHO_E, rhos, sies, gm1s, num, lambdas)

    # -- the rest of the vector surface (eos/eos_base.hpp:422-470,497-501,680-722) ------------
    def _extra(self, fn, in1, in2, out, num, lambdas):
        _require_cuda_f64(in1, in2, out, lambdas)
        n = in1.numel() if num is None else int(num)
        check(fn(self._h, _ptr(in1), _ptr(in2), _ptr(out), n, _ptr(lambdas),
                 self._lambda_stride(lambdas, n), _stream()))
        return out

    def MinInternalEnergyFromDensity(self, rhos, sies, num=None, lambdas=None):
        _require_cuda_f64(rhos, sies, lambdas)
        n = rhos.numel() if num is None else int(num)
        check(self._lib.sgx_MinInternalEnergyFromDensity(
            self._h, _ptr(rhos), _ptr(sies), n, _ptr(lambdas), self._lambda_stride(lambdas, n),
            _stream()))
        return sies

    def InternalEnergyFromDensityPressure(self, rhos, Ps, sies, num=None, lambdas=None):
        return self._extra(self._lib.sgx_InternalEnergyFromDensityPressure, rhos, Ps, sies, num,
                           lambdas)

    def EntropyFromDensityTemperature(self, rhos, temperatures, entropies, num=None, lambdas=None):
        return self._extra(self._lib.sgx_EntropyFromDensityTemperature, rhos, temperatures,
                           entropies, num, lambdas)

    def EntropyFromDensityInternalEnergy(self, rhos, sies, entropies, num=None, lambdas=None):
        return self._extra(self._lib.sgx_EntropyFromDensityInternalEnergy, rhos, sies, entropies,
                           num, lambdas)

    def GibbsFreeEnergyFromDensityTemperature(self, rhos, temperatures, gibbs, num=None,
                                              lambdas=None):
        return self._extra(self._lib.sgx_GibbsFreeEnergyFromDensityTemperature, rhos,
                           temperatures, gibbs, num, lambdas)

    def GibbsFreeEnergyFromDensityInternalEnergy(self, rhos, sies, gibbs, num=None, lambdas=None):
        return self._extra(self._lib.sgx_GibbsFreeEnergyFromDensityInternalEnergy, rhos, sies,
                           gibbs, num, lambdas)

    def DensityEnergyFromPressureTemperature(self, presses, temperatures, rhos, sies, num=None,
                                             lambdas=None):
        """Per cell (eos_base.hpp:630-671 and the models' overloads); rhos is in/out."""
        _require_cuda_f64(presses, temperatures, rhos, sies, lambdas)
        n = presses.numel() if num is None else int(num)
        check(self._lib.sgx_DensityEnergyFromPressureTemperature(
            self._h, _ptr(presses), _ptr(temperatures), _ptr(rhos), _ptr(sies), n, _ptr(lambdas),
            self._lambda_stride(lambdas, n), _stream()))
        return rhos, sies

    def PressureEnergyFromDensityTemperature(self, rhos, temperatures, pressures, sies,
                                             num=None, lambdas=None):
        """Fused P(rho,T)+e(rho,T) (one pass; not in the reference, same results)."""
        _require_cuda_f64(rhos, temperatures, pressures, sies, lambdas)
        n = rhos.numel() if num is None else int(num)
        stride = self._lambda_stride(lambdas, n)
        check(self._lib.sgx_PressureEnergyFromDensityTemperature(
            self._h, _ptr(rhos), _ptr(temperatures), _ptr(pressures), _ptr(sies), n,
            _ptr(lambdas), stride, _stream()))

    def FillEos(self, rhos, temps, energies, presses, cvs, bmods, num, output, lambdas=None):
        """eos/eos_base.hpp:503-531; `output` is a thermalqs bitmask."""
        _require_cuda_f64(rhos, temps, energies, presses, cvs, bmods, lambdas)
        stride = self._lambda_stride(lambdas, int(num))
        check(self._lib.sgx_FillEos(self._h, _ptr(rhos), _ptr(temps), _ptr(energies),
                                    _ptr(presses), _ptr(cvs), _ptr(bmods), int(num),
                                    int(output), _ptr(lambdas), stride, _stream()))

    # -- host-array forms (numpy in / numpy out; staging inside the library) --------------
    def eval_host(self, method, in1, in2, out, lambdas=None):
        import numpy as np

        for a in (in1, in2, out):
            assert isinstance(a, np.ndarray) and a.dtype == np.float64 and a.flags.c_contiguous
        stride = self._lambda_stride(lambdas, in1.size)
        check(self._lib.sgx_eos_eval_host(
            self._h, method, in1.ctypes.data, in2.ctypes.data, out.ctypes.data, in1.size,
            None if lambdas is None else lambdas.ctypes.data, stride))
        return out

    def FillEos_host(self, rhos, temps, energies, presses, cvs, bmods, output, lambdas=None):
        def p(a):
            return None if a is None else a.ctypes.data

        n = next(a for a in (rhos, temps, energies, presses) if a is not None).size
        stride = self._lambda_stride(lambdas, n)
        check(self._lib.sgx_FillEos_host(self._h, p(rhos), p(temps), p(energies), p(presses),
                                         p(cvs), p(bmods), n, int(output), p(lambdas), stride))

    # -- inspection -----------------------------------------------------------------------
    def get_table(self, name):
        import numpy as np

        n = self._lib.sgx_eos_get_table(self._h, name.encode(), None, 0)
        if n < 0:
            raise _lib.SgxError(n, self._lib.sgx_last_error().decode())
        out = np.empty(n, dtype=np.float64)
        m = self._lib.sgx_eos_get_table(self._h, name.encode(), out.ctypes.data, n)
        if m < 0:
            raise _lib.SgxError(m, self._lib.sgx_last_error().decode())
        return out

    def get_scalar(self, name):
        v = C.c_double()
        check(self._lib.sgx_eos_get_scalar(self._h, name.encode(), C.byref(v)))
        return v.value


def _new(fn, *args):
    h = C.c_void_p()
    check(fn(*args, C.byref(h)))
    return h


class IdealGas(EosBase):
    def __init__(self, gm1, Cv):
        lib = _lib.load()
        super().__init__(_new(lib.sgx_eos_create_ideal, gm1, Cv))


class Gruneisen(EosBase):
    def __init__(self, C0, s1, s2, s3, G0, b, rho0, T0, P0, Cv, rho_max=-1.0):
        lib = _lib.load()
        super().__init__(_new(lib.sgx_eos_create_gruneisen, C0, s1, s2, s3, G0, b, rho0, T0,
                              P0, Cv, rho_max))


class DavisReactants(EosBase):
    def __init__(self, rho0, e0, P0, T0, A, B, C_, G0, Z, alpha, Cv0):
        lib = _lib.load()
        super().__init__(_new(lib.sgx_eos_create_davis_reactants, rho0, e0, P0, T0, A, B, C_,
                              G0, Z, alpha, Cv0))


class DavisProducts(EosBase):
    def __init__(self, a, b, k, n, vc, pc, Cv):
        lib = _lib.load()
        super().__init__(_new(lib.sgx_eos_create_davis_products, a, b, k, n, vc, pc, Cv))


class SpinerEOSDependsRhoT(EosBase):
    """Constructed from raw .sp5 contents (tablegen.Sp5Raw) instead of a file name."""

    def root_stats(self, reset=False):
        """(cells through TemperatureFromDensityInternalEnergy, cells among them that needed the
        reference's regula-falsi iterate) since creation / the last reset."""
        out = (C.c_uint64 * 2)()
        check(self._lib.sgx_eos_root_stats(self._h, out, int(reset)))
        return int(out[0]), int(out[1])

    def __init__(self, raw, reproducibility_mode=False, pmin_vapor_dome=False):
        lib = _lib.load()
        self._raw = raw
        super().__init__(_new(lib.sgx_eos_create_spiner_rhot, raw.byref(),
                              int(reproducibility_mode), int(pmin_vapor_dome)))


class SpinerEOSDependsRhoSie(EosBase):
    def __init__(self, raw, reproducibility_mode=False, pmin_vapor_dome=False):
        lib = _lib.load()
        self._raw = raw
        super().__init__(_new(lib.sgx_eos_create_spiner_rhosie, raw.byref(),
                              int(reproducibility_mode), int(pmin_vapor_dome)))


class Helmholtz(EosBase):
    def __init__(self, filename, rad=True, gas=True, coul=True, ion=True, ele=True,
                 newton_raphson=True, lambda_has_lT_slot=True):
        lib = _lib.load()
        opt = _lib.HelmOptions(int(rad), int(gas), int(coul), int(ion), int(ele),
                               int(newton_raphson), int(lambda_has_lT_slot))
        super().__init__(_new(lib.sgx_eos_create_helmholtz, filename.encode(), C.byref(opt)))


# -- modifiers (eos/modifiers/shifted_eos.hpp, scaled_eos.hpp, ramps_eos.hpp) ------------------
# The reference writes ShiftedEOS<T>(T&&, shift) etc.; here the wrapper takes a host-mirror
# object (plain or already modified) and returns a new one that keeps the base alive.  Nesting
# order is the default variant's: BilinearRampEOS(ScaledEOS(ShiftedEOS(T))), each layer optional.
class _Modified(EosBase):
    def __init__(self, handle, base):
        super().__init__(handle)
        self.base = base

    def base_model(self):
        """The unmodified model under every layer."""
        b = self.base
        while isinstance(b, _Modified):
            b = b.base
        return b


def ShiftedEOS(eos, shift):
    return _Modified(_new(_lib.load().sgx_eos_modify_shifted, eos.handle, float(shift)), eos)


def ScaledEOS(eos, scale):
    return _Modified(_new(_lib.load().sgx_eos_modify_scaled, eos.handle, float(scale)), eos)


def BilinearRampEOS(eos, r0, a, b, c):
    return _Modified(_new(_lib.load().sgx_eos_modify_bilinear_ramp, eos.handle, float(r0),
                          float(a), float(b), float(c)), eos)

