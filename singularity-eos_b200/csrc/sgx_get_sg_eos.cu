// get_sg_eos host entry points (the kernel: sgx_get_sg_eos.cuh; its three input modes are
// instantiated in sgx_get_sg_eos_rho_e.cu, _rho_t.cu, _p_t.cu).
#include <cstdlib>
#include <cstring>

#include "sgx_get_sg_eos.cuh"

namespace sgx {
int launch_pte_tab_rhot_sm_mod(const EosSet &set, const MixParamsDev &prm, double *rho,
                               double *vfrac, double *sie, double *temp, double *press,
                               const double *sie_tot, const double *vfrac_tot,
                               const double *Tguess, uint64_t *status, double *residual,
                               uint64_t *stats, int64_t ncell, cudaStream_t s, double *lambda,
                               const unsigned char *skip);
}

using namespace sgx;
using namespace sgx::sg;


extern "C" {

int sgx_get_sg_eos_device(int nmat, int ncell, int cell_dim, int input_int,
                          const int *eos_offsets, const sgx_eos_t *eos, const int *offsets,
                          double *press, double *pmax, double *vol, double *spvol, double *sie,
                          double *temp, double *bmod, double *dpde, double *cv,
                          double *frac_mass, double *frac_vol, double *frac_ie,
                          double *frac_bmod, double *frac_dpde, double *frac_cv,
                          double mass_frac_cutoff, int *nfail, sgx_stream_t stream) {
  DeviceGuard guard;
  if (input_int == -2)
    return fail(SGX_ERR_UNSUPPORTED, "get_sg_eos: the (rho, P) input condition runs "
                "PTESolverFixedP with iterate_t_guess = true in the reference, whose result is "
                "undefined (uninitialised uscale); not built");
  if (input_int != 0 && input_int != 1 && input_int != -3 && input_int != -1)
    return fail(SGX_ERR_INVALID, "get_sg_eos: unknown input_int %d", input_int);
  if (nmat < 1 || nmat > kMaxMat)
    return fail(SGX_ERR_UNSUPPORTED, "get_sg_eos: nmat = %d, this build holds 1..%d", nmat, kMaxMat);
  if (ncell < 0 || cell_dim < ncell) return fail(SGX_ERR_INVALID, "get_sg_eos: bad sizes");
  if (!eos_offsets || !eos || !offsets || !press || !pmax || !vol || !spvol || !sie || !temp ||
      !bmod || !dpde || !cv || !frac_mass || !frac_vol || !frac_ie)
    return fail(SGX_ERR_INVALID, "get_sg_eos: null array");
  if (ncell == 0) return SGX_OK;
  int max_off = 0;
  for (int m = 0; m < nmat; ++m)
    max_off = eos_offsets[m] > max_off ? eos_offsets[m] : max_off;
  if (max_off > kMaxMat)
    return fail(SGX_ERR_UNSUPPORTED, "get_sg_eos: %d EOS objects, this build holds %d", max_off,
                kMaxMat);
  EosSet set;
  set.nmat = max_off;
  SgArgs A;
  for (int m = 0; m < nmat; ++m) {
    const int k = eos_offsets[m] - 1;
    if (k < 0 || !eos[k]) return fail(SGX_ERR_INVALID, "get_sg_eos: bad eos_offsets[%d]", m);
    A.eos_idx[m] = k;
  }
  int device = -1;
  for (int k = 0; k < max_off; ++k) {
    if (!eos[k]) { // an index no material refers to
      set.e[k].kind = SGX_EOS_IDEAL;
      set.e[k].ideal = IdealGas{1.0, 1.0};
      set.mod[k] = ModParams();
      continue;
    }
    if (device < 0) device = eos[k]->device;
    if (eos[k]->device != device) return fail(SGX_ERR_INVALID, "EOS %d lives on another device", k);
    if (!set_material(eos[k], set, k))
      return fail(SGX_ERR_UNSUPPORTED, "get_sg_eos: EOS %d is kind %d", k, eos[k]->kind);
  }
  SGX_CUDA_OK(cudaSetDevice(device));
  sgx_mix_params dflt;
  sgx_mix_params_default(&dflt);
  const MixParamsDev prm = mix_params_to_dev(&dflt);
  A.nmat = nmat;
  A.ncell = ncell;
  A.cell_dim = cell_dim;
  A.offsets = offsets;
  A.press = press, A.pmax = pmax, A.vol = vol, A.spvol = spvol, A.sie = sie, A.temp = temp;
  A.bmod = bmod, A.dpde = dpde, A.cv = cv;
  A.frac_mass = frac_mass, A.frac_vol = frac_vol, A.frac_ie = frac_ie;
  A.frac_bmod = frac_bmod, A.frac_dpde = frac_dpde, A.frac_cv = frac_cv;
  A.mass_frac_cutoff = mass_frac_cutoff;
  A.nfail = nfail;
  const LaunchCfg c = stream_cfg(ncell, 128, 8);
  cudaStream_t st = (cudaStream_t)stream;
  if (input_int == kRhoT)
    launch_get_sg_eos_rho_t(set, prm, A, c, st);
  else if (input_int == kPT)
    launch_get_sg_eos_p_t(set, prm, A, c, st);
  else {
    // (rho, e): cells in which every material participates -- the usual mixed cell -- go through
    // the lane-refill state machine; SGX_SG_EOS_KERNEL=percell solves every cell in the one
    // thread-per-cell kernel as before (same results).
    const char *v = getenv("SGX_SG_EOS_KERNEL");
    const bool staged = nmat >= 2 && nmat <= 4 && !(v && strcmp(v, "percell") == 0);
    char *scratch = nullptr;
    if (staged) {
      const size_t nc = (size_t)ncell, nf = nc * (size_t)nmat;
      // 5 state arrays + lambda caches [ncell][nmat][3] + sie_tot, vfrac_tot, zero Tguess,
      // status words, skip bytes
      const size_t doubles = 5 * nf + 3 * nf + 3 * nc;
      const size_t bytes = doubles * sizeof(double) + nc * sizeof(unsigned long long) +
                           ((nc + 255) & ~size_t(255));
      SGX_CUDA_OK(scratch_alloc_async(reinterpret_cast<void **>(&scratch), bytes, st));
      double *p = reinterpret_cast<double *>(scratch);
      double *s_rho = p, *s_vf = p + nf, *s_sie = p + 2 * nf, *s_temp = p + 3 * nf,
             *s_press = p + 4 * nf, *lam = p + 5 * nf, *sie_tot = p + 8 * nf,
             *vfrac_tot = p + 8 * nf + nc, *Tzero = p + 8 * nf + 2 * nc;
      unsigned long long *status = reinterpret_cast<unsigned long long *>(p + 8 * nf + 3 * nc);
      unsigned char *skip = reinterpret_cast<unsigned char *>(status + nc);
      cudaError_t err = cudaMemsetAsync(lam, 0, 3 * nf * sizeof(double), st); // zeroed caches
      if (err == cudaSuccess) err = cudaMemsetAsync(Tzero, 0, nc * sizeof(double), st);
      if (err != cudaSuccess) {
        cudaFreeAsync(scratch, st);
        return fail(SGX_ERR_CUDA, "get_sg_eos: %s", cudaGetErrorString(err));
      }
      const LaunchCfg cp = stream_cfg(ncell, 256, 8);
      get_sg_eos_prep_kernel<<<cp.blocks, cp.threads, 0, st>>>(A, skip, s_rho, s_vf, s_sie, s_temp,
                                                               s_press, sie_tot, vfrac_tot);
      // the solver's materials in participation order = the caller's material order
      EosSet set2;
      set2.nmat = nmat;
      for (int m = 0; m < nmat; ++m) {
        set2.e[m] = set.e[A.eos_idx[m]];
        set2.mod[m] = set.mod[A.eos_idx[m]];
      }
      // Tguess = 0: PTESolverRhoT is constructed without a guess there (get_sg_eos_rho_e.cpp:
      // 61-64), so GetTguess starts from default_tguess
      const int rc = launch_pte_tab_rhot_sm_mod(
          set2, prm, s_rho, s_vf, s_sie, s_temp, s_press, sie_tot, vfrac_tot, Tzero,
          reinterpret_cast<uint64_t *>(status), nullptr, nullptr, ncell, st, lam, skip);
      if (rc != SGX_OK) {
        cudaFreeAsync(scratch, st);
        return rc;
      }
      A.pre_skip = skip;
      A.pre_rho = s_rho, A.pre_vfrac = s_vf, A.pre_sie = s_sie, A.pre_temp = s_temp;
      A.pre_press = s_press;
      A.pre_status = status;
      count_launch(1);
    }
    launch_get_sg_eos_rho_e(set, prm, A, c, st);
    if (scratch) cudaFreeAsync(scratch, st);
  }
  count_launch();
  SGX_CUDA_OK(cudaGetLastError());
  return SGX_OK;
}

// The reference's signature and memory contract: every array is a HOST array; inputs are
// mirrored to the device, results copied back exactly as get_sg_eos.cpp:236-278 does.
int sgx_get_sg_eos(int nmat, int ncell, int cell_dim, int input_int, int *eos_offsets,
                   const sgx_eos_t *eos, int *offsets, double *press, double *pmax, double *vol,
                   double *spvol, double *sie, double *temp, double *bmod, double *dpde,
                   double *cv, double *frac_mass, double *frac_vol, double *frac_ie,
                   double *frac_bmod, double *frac_dpde, double *frac_cv,
                   double mass_frac_cutoff) {
  DeviceGuard guard;
  if (input_int == -2)
    return fail(SGX_ERR_UNSUPPORTED, "get_sg_eos: the (rho, P) input condition runs "
                "PTESolverFixedP with iterate_t_guess = true in the reference, whose result is "
                "undefined (uninitialised uscale); not built");
  if (input_int != 0 && input_int != 1 && input_int != -3 && input_int != -1)
    return fail(SGX_ERR_INVALID, "get_sg_eos: unknown input_int %d", input_int);
  if (!eos || !eos_offsets || nmat < 1 || !eos[eos_offsets[0] - 1])
    return fail(SGX_ERR_INVALID, "get_sg_eos: null EOS");
  if (ncell <= 0) return ncell == 0 ? SGX_OK : fail(SGX_ERR_INVALID, "negative ncell");
  SGX_CUDA_OK(cudaSetDevice(eos[eos_offsets[0] - 1]->device));
  const bool do_bmod = frac_bmod != nullptr, do_dpde = frac_dpde != nullptr;
  const bool do_cv = frac_dpde != nullptr; // sic: get_sg_eos.cpp:66 tests frac_dpde for frac_cv
  const size_t nc = (size_t)cell_dim, nf = (size_t)cell_dim * nmat;
  // one device allocation: 9 per-cell + 6 per-material arrays, offsets, failure counter
  const size_t doubles = 9 * nc + 6 * nf;
  char *d = nullptr;
  const size_t bytes = doubles * sizeof(double) + (size_t)ncell * sizeof(int) + 256;
  SGX_CUDA_OK(cudaMalloc(&d, bytes));
  double *p = reinterpret_cast<double *>(d);
  double *dv[9], *df[6];
  double *hv[9] = {press, pmax, vol, spvol, sie, temp, bmod, dpde, cv};
  double *hf[6] = {frac_mass, frac_vol, frac_ie, frac_bmod, frac_dpde, frac_cv};
  for (int k = 0; k < 9; ++k, p += nc)
    dv[k] = p;
  for (int k = 0; k < 6; ++k, p += nf)
    df[k] = p;
  int *d_offsets = reinterpret_cast<int *>(p);
  int *d_nfail = d_offsets + ncell + ((ncell & 1) ? 1 : 0);
  cudaError_t err = cudaSuccess;
  auto up = [&](void *dst, const void *src, size_t n) {
    if (err == cudaSuccess && src) err = cudaMemcpy(dst, src, n, cudaMemcpyHostToDevice);
  };
  for (int k = 0; k < 9; ++k)
    up(dv[k], hv[k], nc * sizeof(double));
  for (int k = 0; k < 6; ++k)
    up(df[k], hf[k], nf * sizeof(double));
  up(d_offsets, offsets, (size_t)ncell * sizeof(int));
  if (err == cudaSuccess) err = cudaMemset(d_nfail, 0, sizeof(int));
  int rc = SGX_OK;
  if (err == cudaSuccess) {
    rc = sgx_get_sg_eos_device(nmat, ncell, cell_dim, input_int, eos_offsets, eos, d_offsets,
                               dv[0], dv[1], dv[2], dv[3], dv[4], dv[5], dv[6], dv[7], dv[8],
                               df[0], df[1], df[2], do_bmod ? df[3] : nullptr,
                               do_dpde ? df[4] : nullptr, do_cv ? df[5] : nullptr,
                               mass_frac_cutoff, d_nfail, nullptr);
    if (rc == SGX_OK) err = cudaDeviceSynchronize();
  }
  auto down = [&](void *dst, const void *src, size_t n) {
    if (err == cudaSuccess && rc == SGX_OK && dst)
      err = cudaMemcpy(dst, src, n, cudaMemcpyDeviceToHost);
  };
  // get_sg_eos.cpp:236-262: a cell quantity is copied back unless it is an input of the mode
  const bool r_inp = input_int != kPT, t_inp = input_int == kRhoT || input_int == kPT,
             p_inp = input_int == kPT, s_inp = input_int == 0 || input_int == 1;
  if (!p_inp) down(press, dv[0], nc * sizeof(double));
  down(pmax, dv[1], nc * sizeof(double));
  down(vol, dv[2], nc * sizeof(double));
  if (!r_inp) down(spvol, dv[3], nc * sizeof(double));
  if (!s_inp) down(sie, dv[4], nc * sizeof(double));
  if (!t_inp) down(temp, dv[5], nc * sizeof(double));
  down(bmod, dv[6], nc * sizeof(double));
  down(dpde, dv[7], nc * sizeof(double));
  down(cv, dv[8], nc * sizeof(double));
  down(frac_vol, df[1], nf * sizeof(double));
  down(frac_ie, df[2], nf * sizeof(double));
  if (do_bmod) down(frac_bmod, df[3], nf * sizeof(double));
  if (do_dpde) down(frac_dpde, df[4], nf * sizeof(double));
  if (do_cv) down(frac_cv, df[5], nf * sizeof(double));
  cudaFree(d);
  if (rc != SGX_OK) return rc;
  if (err != cudaSuccess) return fail(SGX_ERR_CUDA, "get_sg_eos: %s", cudaGetErrorString(err));
  return SGX_OK; // like the reference, whose failure count is never copied back (:237-240)
}

} // extern "C"
