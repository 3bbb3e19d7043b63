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

namespace singularity {
using EOS = Variant<IdealGas, Gruneisen, DavisReactants, DavisProducts, Vinet,
                    SpinerEOSDependsRhoT, SpinerEOSDependsRhoSie, Helmholtz>;
} // namespace singularity

#endif
