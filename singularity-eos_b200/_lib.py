"""ctypes binding of libsgx_b200.so -- the C ABI declared in include/sgx_b200.h.

This module only *loads* the CUDA library and declares prototypes; it contains no
numerical code and no CPU fallback.  If the shared object is missing the import of the
compute entry points fails loudly (SgxLibraryMissing).
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "lib", "libsgx_b200.so")


class SgxLibraryMissing(RuntimeError):
    pass


class SgxError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"sgx status {code}: {msg}")
        self.code = code


# ---- POD layouts (must match include/sgx_b200.h) -------------------------------------
class Grid1D(C.Structure):
    _fields_ = [("ngrids", C.c_int), ("npoints", C.c_int * 3), ("min", C.c_double * 3),
                ("max", C.c_double * 3), ("dx", C.c_double * 3)]


_DP = C.POINTER(C.c_double)
SP5_TABLE_NAMES = ["P", "sie", "bMod", "dPdRho", "dPdE", "dTdRho", "dTdE", "dEdRho", "dEdT",
                   "PCold", "sieCold", "bModCold", "dPdRhoCold",
                   "T_re", "P_re", "bMod_re", "dPdRho_re", "dPdE_re", "dTdRho_re", "dTdE_re",
                   "dEdRho_re"]


class Sp5Tables(C.Structure):
    _fields_ = ([("lrho", Grid1D), ("lT", Grid1D), ("lE", Grid1D),
                 ("rho_offset", C.c_double), ("T_offset", C.c_double),
                 ("sie_offset", C.c_double), ("normal_density", C.c_double),
                 ("mean_atomic_mass", C.c_double), ("mean_atomic_number", C.c_double)]
                + [(n, _DP) for n in SP5_TABLE_NAMES])


class HelmOptions(C.Structure):
    _fields_ = [("enable_rad", C.c_int), ("enable_gas", C.c_int), ("enable_coulomb", C.c_int),
                ("gas_ionized", C.c_int), ("gas_degenerate", C.c_int),
                ("use_newton_raphson", C.c_int), ("lambda_has_lT_slot", C.c_int)]


class MixParams(C.Structure):
    """closure/mixed_cell_models.hpp:43-78 (field order of sgx_mix_params)."""
    _fields_ = [("verbose", C.c_int), ("iterate_t_guess", C.c_int),
                ("derivative_eps", C.c_double),
                ("pte_rel_tolerance_p", C.c_double), ("pte_rel_tolerance_t", C.c_double),
                ("pte_rel_tolerance_e", C.c_double), ("pte_rel_tolerance_v", C.c_double),
                ("pte_abs_tolerance_p", C.c_double), ("pte_abs_tolerance_t", C.c_double),
                ("pte_slow_convergence_thresh", C.c_double),
                ("pte_rel_tolerance_p_sufficient", C.c_double),
                ("pte_rel_tolerance_t_sufficient", C.c_double),
                ("pte_rel_tolerance_e_sufficient", C.c_double),
                ("pte_rel_tolerance_v_sufficient", C.c_double),
                ("pte_abs_tolerance_p_sufficient", C.c_double),
                ("pte_abs_tolerance_t_sufficient", C.c_double),
                ("pte_max_iter_per_mat", C.c_uint64),
                ("line_search_alpha", C.c_double),
                ("line_search_max_iter", C.c_uint64),
                ("line_search_fac", C.c_double), ("vfrac_safety_fac", C.c_double),
                ("temperature_limit", C.c_double), ("default_tguess", C.c_double),
                ("min_dtde", C.c_double),
                ("pte_small_step_tries", C.c_uint64),
                ("pte_small_step_thresh", C.c_double), ("pte_max_dpdv", C.c_double),
                ("pte_reduced_system_exclude_idx", C.c_int64)]


# enums
(EOS_IDEAL, EOS_GRUNEISEN, EOS_DAVIS_REACTANTS, EOS_DAVIS_PRODUCTS, EOS_SPINER_RHOT,
 EOS_SPINER_RHOSIE, EOS_HELMHOLTZ, EOS_MODIFIED) = range(8)
(P_FROM_RHO_T, E_FROM_RHO_T, T_FROM_RHO_E, P_FROM_RHO_E, CV_FROM_RHO_T, CV_FROM_RHO_E,
 BMOD_FROM_RHO_T, BMOD_FROM_RHO_E, GAMMA_FROM_RHO_T, GAMMA_FROM_RHO_E) = range(10)
METHOD_NAMES = {
    P_FROM_RHO_T: "PressureFromDensityTemperature",
    E_FROM_RHO_T: "InternalEnergyFromDensityTemperature",
    T_FROM_RHO_E: "TemperatureFromDensityInternalEnergy",
    P_FROM_RHO_E: "PressureFromDensityInternalEnergy",
    CV_FROM_RHO_T: "SpecificHeatFromDensityTemperature",
    CV_FROM_RHO_E: "SpecificHeatFromDensityInternalEnergy",
    BMOD_FROM_RHO_T: "BulkModulusFromDensityTemperature",
    BMOD_FROM_RHO_E: "BulkModulusFromDensityInternalEnergy",
    GAMMA_FROM_RHO_T: "GruneisenParamFromDensityTemperature",
    GAMMA_FROM_RHO_E: "GruneisenParamFromDensityInternalEnergy",
}
PTE_RHOT, PTE_PT, PTE_FIXED_T, PTE_RHOU, PTE_FIXED_P = 0, 1, 2, 3, 4
PTE_NSTATS = 64


class thermalqs:  # base/constants.hpp:23-36
    none = 0
    density = 1 << 0
    specific_internal_energy = 1 << 1
    pressure = 1 << 2
    temperature = 1 << 3
    specific_heat = 1 << 4
    bulk_modulus = 1 << 5


_VP = C.c_void_p
_I64 = C.c_int64
_D = C.c_double

# name -> (restype, argtypes); every symbol include/sgx_b200.h declares
PROTOTYPES = {
    "sgx_abi_version": (C.c_int, []),
    "sgx_last_error": (C.c_char_p, []),
    "sgx_device_count": (C.c_int, []),
    "sgx_set_device": (C.c_int, [C.c_int]),
    "sgx_launch_count": (C.c_uint64, []),
    "sgx_eos_create_ideal": (C.c_int, [_D, _D, C.POINTER(_VP)]),
    "sgx_eos_modify_shifted": (C.c_int, [_VP, _D, C.POINTER(_VP)]),
    "sgx_eos_modify_scaled": (C.c_int, [_VP, _D, C.POINTER(_VP)]),
    "sgx_eos_modify_bilinear_ramp": (C.c_int, [_VP, _D, _D, _D, _D, C.POINTER(_VP)]),
    "sgx_eos_create_gruneisen": (C.c_int, [_D] * 11 + [C.POINTER(_VP)]),
    "sgx_eos_create_davis_reactants": (C.c_int, [_D] * 11 + [C.POINTER(_VP)]),
    "sgx_eos_create_davis_products": (C.c_int, [_D] * 7 + [C.POINTER(_VP)]),
    "sgx_eos_create_spiner_rhot": (C.c_int, [C.POINTER(Sp5Tables), C.c_int, C.c_int,
                                             C.POINTER(_VP)]),
    "sgx_eos_create_spiner_rhosie": (C.c_int, [C.POINTER(Sp5Tables), C.c_int, C.c_int,
                                               C.POINTER(_VP)]),
    "sgx_helm_default_options": (None, [C.POINTER(HelmOptions)]),
    "sgx_eos_create_helmholtz": (C.c_int, [C.c_char_p, C.POINTER(HelmOptions),
                                           C.POINTER(_VP)]),
    "sgx_eos_destroy": (C.c_int, [_VP]),
    "sgx_eos_type": (C.c_int, [_VP]),
    "sgx_eos_nlambda": (C.c_int, [_VP]),
    "sgx_eos_preferred_input": (C.c_ulong, [_VP]),
    "sgx_eos_scratch_bytes": (C.c_uint64, [_VP, C.c_int, _I64]),
    "sgx_eos_maximum_density": (_D, [_VP]),
    "sgx_eos_get_table": (_I64, [_VP, C.c_char_p, _VP, _I64]),
    "sgx_eos_get_scalar": (C.c_int, [_VP, C.c_char_p, C.POINTER(_D)]),
    "sgx_eos_eval": (C.c_int, [_VP, C.c_int, _VP, _VP, _VP, _I64, _VP, C.c_int, _VP]),
    "sgx_PressureFromDensityTemperature": (C.c_int, [_VP, _VP, _VP, _VP, _I64, _VP, C.c_int,
                                                     _VP]),
    "sgx_InternalEnergyFromDensityTemperature": (C.c_int, [_VP, _VP, _VP, _VP, _I64, _VP,
                                                           C.c_int, _VP]),
    "sgx_TemperatureFromDensityInternalEnergy": (C.c_int, [_VP, _VP, _VP, _VP, _I64, _VP,
                                                           C.c_int, _VP]),
    "sgx_PressureFromDensityInternalEnergy": (C.c_int, [_VP, _VP, _VP, _VP, _I64, _VP,
                                                        C.c_int, _VP]),
    "sgx_noalloc_max_cells": (C.c_int64, []),
    "sgx_TemperatureFromDensityInternalEnergy_noalloc": (C.c_int, [_VP, _VP, _VP, _VP, _I64, _VP,
                                                                   C.c_int, _VP]),
    "sgx_PressureEnergyFromDensityTemperature": (C.c_int, [_VP, _VP, _VP, _VP, _VP, _I64, _VP,
                                                           C.c_int, _VP]),
    "sgx_FillEos": (C.c_int, [_VP] * 7 + [_I64, C.c_ulong, _VP, C.c_int, _VP]),
    "sgx_eos_eval_host": (C.c_int, [_VP, C.c_int, _VP, _VP, _VP, _I64, _VP, C.c_int]),
    "sgx_FillEos_host": (C.c_int, [_VP] * 7 + [_I64, C.c_ulong, _VP, C.c_int]),
    "sgx_mix_params_default": (None, [C.POINTER(MixParams)]),
    "sgx_pte_solve": (C.c_int, [C.c_int, C.POINTER(_VP), C.c_int, _I64] + [_VP] * 8 +
                      [C.POINTER(MixParams), _VP, _VP, _VP, _VP]),
    "sgx_pte_solve_cached": (C.c_int, [C.c_int, C.POINTER(_VP), C.c_int, _I64] + [_VP] * 8 +
                      [C.POINTER(MixParams), _VP, _VP, _VP, _VP, _VP]),
    "sgx_pte_required_scratch": (C.c_int, [C.c_int, C.c_int]),
    "sgx_pte_solve_host": (C.c_int, [C.c_int, C.POINTER(_VP), C.c_int, _I64] + [_VP] * 8 +
                           [C.POINTER(MixParams), _VP, _VP]),
    "sgx_fastmath_eval": (C.c_int, [C.c_int, _VP, _VP, _I64, _VP]),
    "sgx_libm_eval": (C.c_int, [C.c_int, _VP, _VP, _VP, _I64, _VP]),
    "sgx_eos_root_stats": (C.c_int, [_VP, C.POINTER(C.c_uint64), C.c_int]),
    "sgx_eos_retain": (C.c_int, [_VP]),
    "sgx_eos_serialized_size": (C.c_uint64, [_VP]),
    "sgx_eos_serialize": (C.c_int, [_VP, _VP, C.c_uint64]),
    "sgx_eos_deserialize": (C.c_int, [_VP, C.c_uint64, C.POINTER(_VP)]),
    "sgx_get_sg_eos": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, _VP, C.POINTER(_VP), _VP] +
                       [_VP] * 15 + [C.c_double]),
    "sgx_get_sg_eos_device": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, _VP, C.POINTER(_VP),
                                         _VP] + [_VP] * 15 + [C.c_double, _VP, _VP]),
    "sgx_MinInternalEnergyFromDensity": (C.c_int, [_VP, _VP, _VP, _I64, _VP, C.c_int, _VP]),
    "sgx_InternalEnergyFromDensityPressure": (C.c_int, [_VP, _VP, _VP, _VP, _I64, _VP, C.c_int, _VP]),
    "sgx_EntropyFromDensityTemperature": (C.c_int, [_VP, _VP, _VP, _VP, _I64, _VP, C.c_int, _VP]),
    "sgx_EntropyFromDensityInternalEnergy": (C.c_int, [_VP, _VP, _VP, _VP, _I64, _VP, C.c_int, _VP]),
    "sgx_GibbsFreeEnergyFromDensityTemperature": (C.c_int, [_VP, _VP, _VP, _VP, _I64, _VP, C.c_int, _VP]),
    "sgx_GibbsFreeEnergyFromDensityInternalEnergy": (C.c_int, [_VP, _VP, _VP, _VP, _I64, _VP, C.c_int, _VP]),
    "sgx_DensityEnergyFromPressureTemperature": (C.c_int, [_VP, _VP, _VP, _VP, _VP, _I64, _VP, C.c_int, _VP]),
    "sgx_solve_Ax_b": (C.c_int, [C.c_int, _VP, _VP, _VP, _I64, _VP]),
    "sgx_comm_unique_id": (C.c_int, [C.c_char_p]),
    "sgx_comm_create": (C.c_int, [C.c_char_p, C.c_int, C.c_int, C.POINTER(_VP)]),
    "sgx_comm_from_nccl": (C.c_int, [_VP, C.POINTER(_VP)]),
    "sgx_comm_destroy": (C.c_int, [_VP]),
    "sgx_stats_allreduce": (C.c_int, [_VP, _VP, C.c_int, _VP]),
    "sgx_microbench_fp64": (C.c_int, [C.c_int, C.POINTER(C.c_double)]),
}

_lib = None


def load():
    """dlopen libsgx_b200.so and declare prototypes.  Never falls back to anything."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise SgxLibraryMissing(
            f"{LIB_PATH} not found: build it with `python __graft_entry__.py build` "
            "(nvcc, sm_100a).  There is no CPU fallback for the sgx_b200 hot path.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError if the .so lacks a declared symbol
        fn.restype = res
        fn.argtypes = args
    if lib.sgx_abi_version() != 1:
        raise SgxLibraryMissing("libsgx_b200.so ABI version mismatch; rebuild")
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        raise SgxError(rc, load().sgx_last_error().decode(errors="replace"))
