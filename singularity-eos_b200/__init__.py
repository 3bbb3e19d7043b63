"""singularity-eos_b200: B200-native (sm_100a) cell-parallel EOS lookups and PTE solves.

Drop-in for the hot path of lanl/singularity-eos behind a C ABI (include/sgx_b200.h).
The directory name contains a hyphen (it follows the repository naming), so import it
through `__graft_entry__.load_package()` or as `singularity_eos_b200` after that call.
"""
from . import _lib, closure, dist, eos, rng, tablegen, workloads  # noqa: F401
from ._lib import SgxError, SgxLibraryMissing, thermalqs  # noqa: F401
from .eos import (BilinearRampEOS, ScaledEOS, ShiftedEOS, DavisProducts, DavisReactants, Gruneisen, Helmholtz, IdealGas,  # noqa: F401
                  SpinerEOSDependsRhoSie, SpinerEOSDependsRhoT)
