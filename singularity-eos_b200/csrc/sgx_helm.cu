// Helmholtz EOS: ASCII table loader, cell-major repack, FillEos / vector-lookup kernels.
// Reference: singularity-eos/eos/eos_helmholtz.hpp (InitDataFile_ :962-989; FillEos :818-865;
// per-quantity vector methods :525-661 all funnel into FillEos).
//
// Kernel design: thread-per-cell, grid-stride, 128-thread blocks.  The work per cell is
// ~1.5 k FP64 operations (five 36-term biquintic sums, seven 16-term bicubic sums, the
// radiation / ion / Coulomb closed forms, ~10 libm calls) against 72 algorithmic bytes, so the
// FP64 pipe -- not HBM -- bounds it.  The packed cell records (18 MiB) stay
// L2-resident; each lookup reads 21 aligned 32-byte sectors with 16-byte loads.
#include <cerrno>
#include <cmath>
#include <cstdlib>
#include <fstream>
#include <sstream>

#include "sgx_internal.hpp"

namespace sgx {

namespace {
// Block layout of helm_table.dat (eos_helmholtz.hpp:962-975): j (temperature) outer,
// i (density) inner; columns in the order the reference passes the DataBoxes.
constexpr int kBlockCols[4] = {9, 4, 4, 4};
enum F9 { F = 0, FD = 1, FT = 2, FDD = 3, FTT = 4, FDT = 5, FDDT = 6, FDTT = 7, FDDTT = 8 };

// order of the 9 free-energy tables inside fi[0..35] (eos_helmholtz.hpp:1027-1062):
// f, ft, ftt, fd, fdd, fdt, fddt, fdtt, fddtt -- four corners each
constexpr int kFiOrder9[9] = {F, FT, FTT, FD, FDD, FDT, FDDT, FDTT, FDDTT};
// order inside the bicubic groups (:1176-1191): X, Xt, Xd, Xdt with file columns X Xd Xt Xdt
constexpr int kFiOrder4[4] = {0, 2, 1, 3};

bool read_all(const char *path, std::vector<double> &vals, size_t expect) {
  std::ifstream f(path, std::ios::binary);
  if (!f.is_open()) return false;
  std::stringstream ss;
  ss << f.rdbuf();
  const std::string s = ss.str();
  vals.clear();
  vals.reserve(expect);
  const char *p = s.c_str();
  char *end = nullptr;
  while (vals.size() < expect) {
    errno = 0;
    const double v = std::strtod(p, &end); // correctly rounded, like operator>>
    if (end == p) break;
    vals.push_back(v);
    p = end;
  }
  return vals.size() == expect;
}
} // namespace

int build_helmholtz(const char *path, const sgx_helm_options *opt, sgx_eos_s *e) {
  using namespace helm;
  if (!path) return fail(SGX_ERR_INVALID, "null Helmholtz table path");
  const size_t nnode = (size_t)NTEMP * NRHO;
  std::vector<double> raw;
  if (!read_all(path, raw, nnode * 21))
    return fail(SGX_ERR_IO, "HelmElectrons file %s not found or short (need %zu numbers)", path,
                nnode * 21);
  return build_helmholtz_raw(raw.data(), raw.size(), opt, e);
}

// From the 21 x 101 x 271 numbers of the table file, in file order (also the deserialization path)
int build_helmholtz_raw(const double *raw, size_t nraw, const sgx_helm_options *opt,
                        sgx_eos_s *e) {
  using namespace helm;
  const size_t nnode = (size_t)NTEMP * NRHO;
  if (!raw || nraw != nnode * 21)
    return fail(SGX_ERR_INVALID, "Helmholtz table: need %zu numbers", nnode * 21);
  sgx_helm_options o;
  if (opt) {
    o = *opt;
  } else {
    sgx_helm_default_options(&o);
  }
  { // the creation recipe, for sgx_eos_serialize
    Recipe r(SGX_EOS_HELMHOLTZ);
    r.put(&o, sizeof(o));
    r.put_array(raw, nraw);
    e->recipe = std::move(r.bytes);
  }
  // node(j, i, block, col)
  size_t block_off[4];
  block_off[0] = 0;
  for (int b = 1; b < 4; ++b)
    block_off[b] = block_off[b - 1] + nnode * kBlockCols[b - 1];
  auto node = [&](int b, int col, int j, int i) {
    return raw[block_off[b] + ((size_t)j * NRHO + i) * kBlockCols[b] + col];
  };

  TableStore &ts = e->tables;
  // RegularGrid1D(min, max, N): dx = (max - min) / (N - 1), idx = 1 / dx
  const double lR_dx = (lRhoMax - lRhoMin) / ((double)(NRHO - 1));
  const double lT_dx = (lTMax - lTMin) / ((double)(NTEMP - 1));
  // rho_(i) = pow10(lRhoRange.x(i)), T_(j) likewise (:976-985); pow10 = exp(ln10 * x)
  std::vector<double> rho(NRHO), T(NTEMP);
  for (int i = 0; i < NRHO; ++i)
    rho[i] = std::exp(LN10 * (i * lR_dx + lRhoMin));
  for (int j = 0; j < NTEMP; ++j)
    T[j] = std::exp(LN10 * (j * lT_dx + lTMin));
  const size_t off_rho = ts.add("rho_", rho.data(), NRHO);
  const size_t off_T = ts.add("T_", T.data(), NTEMP);
  // the raw node tables are kept for inspection (sgx_eos_get_table): "f", "fd", ...
  static const char *names9[9] = {"f", "fd", "ft", "fdd", "ftt", "fdt", "fddt", "fdtt", "fddtt"};
  static const char *names4[3][4] = {{"dpdf", "dpdfd", "dpdft", "dpdfdt"},
                                     {"ef", "efd", "eft", "efdt"},
                                     {"xf", "xfd", "xft", "xfdt"}};
  std::vector<double> tmp(nnode);
  for (int c = 0; c < 9; ++c) {
    for (int j = 0; j < NTEMP; ++j)
      for (int i = 0; i < NRHO; ++i)
        tmp[(size_t)j * NRHO + i] = node(0, c, j, i);
    ts.add(names9[c], tmp.data(), nnode);
  }
  for (int b = 0; b < 3; ++b)
    for (int c = 0; c < 4; ++c) {
      for (int j = 0; j < NTEMP; ++j)
        for (int i = 0; i < NRHO; ++i)
          tmp[(size_t)j * NRHO + i] = node(b + 1, c, j, i);
      ts.add(names4[b][c], tmp.data(), nnode);
    }
  // cell-major records
  const size_t ncell = (size_t)NCELL_T * NCELL_R;
  const size_t off_cells = ts.add("cells", nullptr, ncell * REC);
  double *cells = ts.host.data() + off_cells;
  for (int j = 0; j < NCELL_T; ++j)
    for (int i = 0; i < NCELL_R; ++i) {
      double *r = cells + ((size_t)j * NCELL_R + i) * REC;
      int k = 0;
      for (int t = 0; t < 9; ++t) {
        const int c = kFiOrder9[t];
        r[k++] = node(0, c, j, i);
        r[k++] = node(0, c, j, i + 1);
        r[k++] = node(0, c, j + 1, i);
        r[k++] = node(0, c, j + 1, i + 1);
      }
      for (int b = 0; b < 3; ++b)
        for (int t = 0; t < 4; ++t) {
          const int c = kFiOrder4[t];
          r[k++] = node(b + 1, c, j, i);
          r[k++] = node(b + 1, c, j, i + 1);
          r[k++] = node(b + 1, c, j + 1, i);
          r[k++] = node(b + 1, c, j + 1, i + 1);
        }
    }
  int rc = upload_tables(e);
  if (rc) return rc;
  HelmDev &h = e->helm;
  h.cells = ts.dev + off_cells;
  h.rho_ = ts.dev + off_rho;
  h.T_ = ts.dev + off_T;
  h.lR_min = lRhoMin;
  h.lR_dx = lR_dx;
  h.lR_idx = 1 / lR_dx;
  h.lT_min = lTMin;
  h.lT_dx = lT_dx;
  h.lT_idx = 1 / lT_dx;
  // `const Real log10e = std::log10(M_E)` (:1522): GCC folds this call at compile time to the
  // correctly rounded value 0x3FDBCB7B1526E50E, which is what the reference's binary holds.
  h.log10e = 0.43429448190325182;
  h.enable_rad = o.enable_rad != 0;
  h.enable_gas = o.enable_gas != 0;
  h.enable_coulomb = o.enable_coulomb != 0;
  h.gas_ionized = o.gas_ionized != 0;
  h.gas_degenerate = o.gas_degenerate != 0;
  h.use_newton_raphson = o.use_newton_raphson != 0;
  h.lambda_has_lT_slot = o.lambda_has_lT_slot != 0;
  ts.scalars["lTMin"] = lTMin;
  ts.scalars["lTMax"] = lTMax;
  ts.scalars["lRhoMin"] = lRhoMin;
  ts.scalars["lRhoMax"] = lRhoMax;
  return SGX_OK;
}

// ---------------------------------------------------------------------------------------
// kernels
// ---------------------------------------------------------------------------------------
// FillEos vector overload (eos_base.hpp:503-518) for Helmholtz.  OUTPUT is a compile-time
// mask for the two benchmarked modes (dead outputs and their arithmetic are pruned), 0 =
// run-time mask.
template <unsigned long OUTPUT, int MINB = 1>
__global__ void __launch_bounds__(128, MINB)
    helm_fill_kernel(const __grid_constant__ HelmDev eos, double *__restrict__ rho,
                     double *__restrict__ T, double *__restrict__ en, double *__restrict__ P,
                     double *__restrict__ cv, double *__restrict__ bmod, const int64_t n,
                     const unsigned long output_rt, double *__restrict__ lambda,
                     const int lstride, int *__restrict__ err) {
  const unsigned long output = OUTPUT ? OUTPUT : output_rt;
  const unsigned long input = ~output;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nthr = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = tid; i < n; i += nthr) {
    double r = 0, t = 0, e = 0, p = 0, c = 0, b = 0;
    if ((input & TQ_DENSITY) && rho) r = rho[i];
    if ((input & TQ_TEMPERATURE) && T) t = T[i];
    if ((input & TQ_SIE) && en) e = en[i];
    if ((input & TQ_PRESSURE) && P) p = P[i];
    const bool ok = eos.fill(r, t, e, p, c, b, output, lambda ? lambda + i * lstride : nullptr);
    if (!ok) {
      if (err) *err = 1;
      continue;
    }
    if ((output & TQ_TEMPERATURE) && T) T[i] = t;
    if ((output & TQ_SIE) && en) en[i] = e;
    if ((output & TQ_PRESSURE) && P) P[i] = p;
    if ((output & TQ_CV) && cv) cv[i] = c;
    if ((output & TQ_BMOD) && bmod) bmod[i] = b;
  }
}

// OUT[i] = eos.METHOD(IN1[i], IN2[i], lambda[i]) -- every per-quantity method of the
// reference is FillEos with a one- or two-bit mask (eos_helmholtz.hpp:525-631), except the
// Grueneisen parameter (:632-661).
__global__ void __launch_bounds__(128)
    helm_eval_kernel(const __grid_constant__ HelmDev eos, const int method,
                     const double *__restrict__ in1, const double *__restrict__ in2,
                     double *__restrict__ out, const int64_t n, double *__restrict__ lambda,
                     const int lstride) {
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nthr = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = tid; i < n; i += nthr) {
    double *lam = lambda + i * lstride;
    double r = in1[i], t = 0, e = 0, p = 0, c = 0, b = 0, res = 0;
    switch (method) {
    case SGX_P_FROM_RHO_T:
      t = in2[i];
      eos.fill(r, t, e, p, c, b, TQ_PRESSURE, lam);
      res = p;
      break;
    case SGX_E_FROM_RHO_T:
      t = in2[i];
      eos.fill(r, t, e, p, c, b, TQ_SIE, lam);
      res = e;
      break;
    case SGX_T_FROM_RHO_E:
      e = in2[i];
      eos.fill(r, t, e, p, c, b, TQ_TEMPERATURE, lam);
      res = t;
      break;
    case SGX_P_FROM_RHO_E:
      e = in2[i];
      eos.fill(r, t, e, p, c, b, TQ_PRESSURE | TQ_TEMPERATURE, lam);
      res = p;
      break;
    case SGX_CV_FROM_RHO_T:
      t = in2[i];
      eos.fill(r, t, e, p, c, b, TQ_CV, lam);
      res = c;
      break;
    case SGX_CV_FROM_RHO_E:
      e = in2[i];
      eos.fill(r, t, e, p, c, b, TQ_CV | TQ_TEMPERATURE, lam);
      res = c;
      break;
    case SGX_BMOD_FROM_RHO_T:
      t = in2[i];
      eos.fill(r, t, e, p, c, b, TQ_BMOD, lam);
      res = b;
      break;
    case SGX_BMOD_FROM_RHO_E:
      e = in2[i];
      eos.fill(r, t, e, p, c, b, TQ_BMOD | TQ_TEMPERATURE, lam);
      res = b;
      break;
    case SGX_GAMMA_FROM_RHO_T:
      res = eos.gruneisen(r, in2[i], false, lam);
      break;
    default:
      res = eos.gruneisen(r, in2[i], true, lam);
      break;
    }
    out[i] = res;
  }
}

static int helm_check_lambda(const sgx_eos_s *e, const double *lambda, int stride) {
  // SafeMustGet<MeanAtomicMass/Number> aborts on a null lambda (eos_helmholtz.hpp:830-833)
  if (!lambda)
    return fail(SGX_ERR_INVALID, "Helmholtz needs a lambda with Abar, Zbar (SafeMustGet)");
  const int need = e->helm.lambda_has_lT_slot ? 3 : 2;
  if (stride < need)
    return fail(SGX_ERR_INVALID, "Helmholtz lambda_stride %d < %d", stride, need);
  return SGX_OK;
}

int launch_helm_fill(sgx_eos_s *e, double *rho, double *T, double *en, double *P, double *cv,
                     double *bmod, int64_t n, unsigned long output, double *lambda, int stride,
                     cudaStream_t s) {
  int rc = helm_check_lambda(e, lambda, stride);
  if (rc) return rc;
  if (output & TQ_DENSITY)
    return fail(SGX_ERR_UNSUPPORTED, "Density output not supported by this EOS");
  if ((output & TQ_TEMPERATURE) && (output & TQ_SIE))
    return fail(SGX_ERR_UNSUPPORTED,
                "Either specific internal energy or temperature must be provided.");
  if (!rho) return fail(SGX_ERR_INVALID, "null density array");
  const LaunchCfg c = stream_cfg(n, 128, 8);
  constexpr unsigned long kModeI = TQ_SIE | TQ_PRESSURE | TQ_CV | TQ_BMOD;
  constexpr unsigned long kModeII = TQ_TEMPERATURE | TQ_PRESSURE | TQ_CV | TQ_BMOD;
  // mode (i): 128 registers / 4 blocks per SM measured 6 % ahead of 204 registers / 2 blocks
  if (output == kModeI && T && en && P && cv && bmod) {
    // SGX_HELM_MINB=3|5: the register-budget study of DESIGN.md 5.4 (168 / 96 registers)
    static const int minb = [] {
      const char *v = getenv("SGX_HELM_MINB");
      return v ? atoi(v) : 4;
    }();
    if (minb == 3)
      helm_fill_kernel<kModeI, 3><<<c.blocks, c.threads, 0, s>>>(
          e->helm, rho, T, en, P, cv, bmod, n, output, lambda, stride, nullptr);
    else if (minb == 5)
      helm_fill_kernel<kModeI, 5><<<c.blocks, c.threads, 0, s>>>(
          e->helm, rho, T, en, P, cv, bmod, n, output, lambda, stride, nullptr);
    else
      helm_fill_kernel<kModeI, 4><<<c.blocks, c.threads, 0, s>>>(
          e->helm, rho, T, en, P, cv, bmod, n, output, lambda, stride, nullptr);
  } else if (output == kModeII && T && en && P && cv && bmod) {
    helm_fill_kernel<kModeII><<<c.blocks, c.threads, 0, s>>>(e->helm, rho, T, en, P, cv, bmod,
                                                             n, output, lambda, stride, nullptr);
  } else {
    helm_fill_kernel<0ul><<<c.blocks, c.threads, 0, s>>>(e->helm, rho, T, en, P, cv, bmod, n,
                                                         output, lambda, stride, nullptr);
  }
  count_launch();
  SGX_CUDA_OK(cudaGetLastError());
  return SGX_OK;
}

int launch_helm_eval(sgx_eos_s *e, int method, const double *in1, const double *in2,
                     double *out, int64_t n, double *lambda, int stride, cudaStream_t s) {
  int rc = helm_check_lambda(e, lambda, stride);
  if (rc) return rc;
  if (method < 0 || method >= SGX_METHOD_COUNT)
    return fail(SGX_ERR_INVALID, "unknown method %d", method);
  const LaunchCfg c = stream_cfg(n, 128, 8);
  helm_eval_kernel<<<c.blocks, c.threads, 0, s>>>(e->helm, method, in1, in2, out, n, lambda,
                                                  stride);
  count_launch();
  SGX_CUDA_OK(cudaGetLastError());
  return SGX_OK;
}

} // namespace sgx
