// The variant a host code would build with the plugin on (as example/plugin/dust/
// dust_variant.hpp:130-141 does for Dust): the reference's own analytic / tabulated types next to
// their B200-backed counterparts.
#ifndef SGX_PLUGIN_VARIANT_HPP_
#define SGX_PLUGIN_VARIANT_HPP_

#include <singularity-eos/eos/eos_variant.hpp>

#include "sgx_make.hpp"

namespace singularity {
#ifdef SINGULARITY_USE_SPINER_WITH_HDF5
using SgxEOS = Variant<IdealGas, Gruneisen, DavisReactants, DavisProducts, SpinerEOSDependsRhoT,
                       SpinerEOSDependsRhoSie, SgxIdealGas, SgxGruneisen, SgxDavisReactants,
                       SgxDavisProducts, SgxSpinerRhoT, SgxSpinerRhoSie>;
#else
using SgxEOS = Variant<IdealGas, Gruneisen, DavisReactants, DavisProducts, SgxIdealGas,
                       SgxGruneisen, SgxDavisReactants, SgxDavisProducts>;
#endif
} // namespace singularity
#endif // SGX_PLUGIN_VARIANT_HPP_
