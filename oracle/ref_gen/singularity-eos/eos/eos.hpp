// TEST INFRASTRUCTURE ONLY (oracle build).
// Plays the role of the generated <singularity-eos/eos/eos.hpp>
// (reference: singularity-eos/eos/eos.hpp.in:17 includes SINGULARITY_VARIANT).
// A narrow Variant holding only the models on the hot path, so the
// reference's closure header (closure/mixed_cell_models.hpp:24) finds singularity::EOS.
#ifndef SGX_ORACLE_NARROW_EOS_HPP_
#define SGX_ORACLE_NARROW_EOS_HPP_

#include <ports-of-call/portability.hpp>
#include <singularity-eos/base/constants.hpp>
#include <singularity-eos/base/eos_error.hpp>
#include <singularity-eos/base/variadic_utils.hpp>
#include <singularity-eos/eos/eos_base.hpp>
#include <singularity-eos/eos/eos_variant.hpp>

#include <singularity-eos/eos/eos_davis.hpp>
#include <singularity-eos/eos/eos_gruneisen.hpp>
#include <singularity-eos/eos/eos_helmholtz.hpp>
#include <singularity-eos/eos/eos_ideal.hpp>
#include <singularity-eos/eos/eos_spiner.hpp>
#include <singularity-eos/eos/eos_vinet.hpp>

#include <singularity-eos/eos/modifiers/ramps_eos.hpp>
#include <singularity-eos/eos/modifiers/scaled_eos.hpp>
#include <singularity-eos/eos/modifiers/shifted_eos.hpp>

namespace singularity {
// The seven modifier combinations of the default variant (eos/eos_type_lists.hpp:69-94) around
// three of the models (an analytic one that prefers pressure, one that prefers energy, and the
// tabulated one with a lambda cache); more bases only lengthen the oracle build.
#define SGX_ORACLE_MODS(T)                                                              \
  ShiftedEOS<T>, ScaledEOS<T>, ScaledEOS<ShiftedEOS<T>>, BilinearRampEOS<T>,              \
      BilinearRampEOS<ShiftedEOS<T>>, BilinearRampEOS<ScaledEOS<T>>,                      \
      BilinearRampEOS<ScaledEOS<ShiftedEOS<T>>>
using EOS = Variant<IdealGas, Gruneisen, DavisReactants, DavisProducts, Vinet,
                    SpinerEOSDependsRhoT, SpinerEOSDependsRhoSie, Helmholtz,
                    SGX_ORACLE_MODS(IdealGas), SGX_ORACLE_MODS(Gruneisen),
                    SGX_ORACLE_MODS(SpinerEOSDependsRhoT)>;
#undef SGX_ORACLE_MODS
} // namespace singularity

#endif
