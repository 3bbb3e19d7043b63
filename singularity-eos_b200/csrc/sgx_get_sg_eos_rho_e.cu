// get_sg_eos, input mode kRhoE (sgx_get_sg_eos.cuh): a translation unit of its own so the modes
// compile in parallel.
#include "sgx_get_sg_eos.cuh"

namespace sgx {
void launch_get_sg_eos_rho_e(const EosSet &set, const MixParamsDev &prm, const sg::SgArgs &A,
                             const LaunchCfg &c, cudaStream_t st) {
  sg::launch_get_sg_eos_mode<sg::kRhoE>(set, prm, A, c, st);
}
} // namespace sgx
