// Host-side internals of libsgx_b200 shared by the translation units: the opaque EOS handle,
// error plumbing, launch accounting and pinned staging buffers.  Not part of the public ABI.
#ifndef SGX_INTERNAL_HPP_
#define SGX_INTERNAL_HPP_

#include <cuda_runtime.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <utility>
#include <vector>

#include "../../include/sgx_b200.h"
#include "sgx_helmholtz.cuh"
#include "sgx_pte.cuh"
#include "sgx_spiner.cuh"

namespace sgx {

// thread-local last-error text
inline std::string &last_error() {
  static thread_local std::string e;
  return e;
}
inline int fail(int code, const char *fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  last_error() = buf;
  return code;
}
#define SGX_CUDA_OK(expr)                                                                \
  do {                                                                                   \
    cudaError_t _e = (expr);                                                             \
    if (_e != cudaSuccess)                                                               \
      return sgx::fail(SGX_ERR_CUDA, "%s failed: %s (%s:%d)", #expr,                     \
                       cudaGetErrorString(_e), __FILE__, __LINE__);                      \
  } while (0)

// Entry points switch to the handle's device (tables are per device); the caller's current
// device is put back when the call returns, so a single-process multi-GPU host sees no side effect.
struct DeviceGuard {
  int saved = -1;
  DeviceGuard() {
    if (cudaGetDevice(&saved) != cudaSuccess) saved = -1;
  }
  ~DeviceGuard() {
    int now = -1;
    if (saved >= 0 && cudaGetDevice(&now) == cudaSuccess && now != saved) cudaSetDevice(saved);
  }
  DeviceGuard(const DeviceGuard &) = delete;
  DeviceGuard &operator=(const DeviceGuard &) = delete;
};

// A caller's host range for the duration of one *_host call.  Pageable memory makes
// cudaMemcpyAsync synchronous and staged by the driver, which serialises the chunk pipeline;
// ranges the caller has not pinned are page-locked here (cudaHostRegister) and released when the
// call returns.  SGX_HOST_PIN=none disables it (the copies then run at pageable speed); a failed
// registration (locked-memory limit) silently falls back to the same.
struct HostPin {
  void *ptr = nullptr;
  HostPin() {}
  HostPin(const void *p, size_t bytes);
  ~HostPin() {
    if (ptr) cudaHostUnregister(ptr);
  }
  HostPin(HostPin &&o) noexcept : ptr(o.ptr) { o.ptr = nullptr; }
  HostPin(const HostPin &) = delete;
  HostPin &operator=(const HostPin &) = delete;
};

extern std::atomic<unsigned long long> g_launches;
inline void count_launch(int n = 1) { g_launches.fetch_add(n, std::memory_order_relaxed); }

// Named host copies of every table + where it sits in the single device allocation.
struct TableStore {
  std::vector<double> host;                                          // concatenated, padded
  std::map<std::string, std::pair<size_t, size_t>> index;            // name -> (offset, count)
  std::map<std::string, double> scalars;
  double *dev = nullptr;
  // 256-byte aligned slots so vector loads / TMA-friendly strides stay possible
  size_t add(const std::string &name, const double *src, size_t n) {
    const size_t off = (host.size() + 31) & ~size_t(31);
    host.resize(off + n, 0.0);
    if (src) std::copy(src, src + n, host.begin() + off);
    index[name] = {off, n};
    return off;
  }
  double *h(const std::string &name) { return host.data() + index.at(name).first; }
  const double *d(const std::string &name) const { return dev + index.at(name).first; }
  bool has(const std::string &name) const { return index.count(name) != 0; }
};

// Byte stream of a handle's creation arguments: "SGXEOS01", int32 kind, then records.
struct Recipe {
  std::vector<char> bytes;
  explicit Recipe(int kind) {
    put("SGXEOS01", 8);
    const int32_t k = kind;
    put(&k, sizeof(k));
  }
  void put(const void *p, size_t n) {
    const char *c = static_cast<const char *>(p);
    bytes.insert(bytes.end(), c, c + n);
  }
  void put_array(const double *p, size_t n) { // count (0 = NULL), then the doubles
    const uint64_t cnt = p ? n : 0;
    put(&cnt, sizeof(cnt));
    if (cnt) put(p, cnt * sizeof(double));
  }
};
struct RecipeReader {
  const char *p, *end;
  bool ok = true;
  RecipeReader(const char *src, size_t n) : p(src), end(src + n) {}
  bool get(void *dst, size_t n) {
    if (!ok || (size_t)(end - p) < n) return ok = false;
    std::memcpy(dst, p, n);
    p += n;
    return true;
  }
  // points INTO the stream (8-byte alignment is the caller's to provide) or nullptr
  bool get_array(std::vector<double> &store, const double *&out) {
    uint64_t cnt = 0;
    if (!get(&cnt, sizeof(cnt))) return false;
    if ((uint64_t)(end - p) < cnt * sizeof(double)) return ok = false;
    store.resize(cnt);
    if (cnt) std::memcpy(store.data(), p, cnt * sizeof(double));
    p += cnt * sizeof(double);
    out = cnt ? store.data() : nullptr;
    return true;
  }
};

} // namespace sgx

struct sgx_eos_s {
  std::atomic<int> refs{1}; // sgx_eos_retain / sgx_eos_destroy
  int kind = -1;
  int device = 0;
  sgx::AnyEos analytic;    // kinds 0-3
  sgx::SpinerRhoT rhot;    // kind 4
  sgx::SpinerRhoSie rhosie; // kind 5
  sgx::HelmDev helm;       // kind 6
  sgx::TableStore tables;
  // pinned staging for the *_host entry points (grown on demand)
  void *pinned = nullptr;
  size_t pinned_bytes = 0;
  void *dscratch = nullptr;
  size_t dscratch_bytes = 0;
  cudaStream_t own_stream = nullptr;
  cudaStream_t host_streams[4] = {nullptr, nullptr, nullptr, nullptr}; // chunk pipeline
  // SpinerRhoT T(rho,sie): device uint64[2] = {cells seen, cells sent to the reference iterate}
  void *root_stats = nullptr;
  // device copy of `rhot` / `rhosie` (kinds 4, 5): what a mixed-cell EosSet points at
  void *dev_struct = nullptr;
  // Everything the handle was created from (parameters / raw tables / options), as the byte
  // stream sgx_eos_serialize hands out and sgx_eos_deserialize rebuilds a handle from.
  std::vector<char> recipe;
  // kind SGX_EOS_MODIFIED: the unmodified model (retained) and the layers around it
  sgx_eos_s *base = nullptr;
  sgx::ModParams mod;
};

namespace sgx {
// One material of a mixed-cell EosSet from a handle (modified handles: base model + layers).
// Returns false for models that cannot sit in a mixed cell (Helmholtz).
inline bool set_material(const sgx_eos_s *h, EosSet &set, const int k) {
  const sgx_eos_s *b = h->kind == SGX_EOS_MODIFIED ? h->base : h;
  set.mod[k] = h->kind == SGX_EOS_MODIFIED ? h->mod : ModParams();
  if (b->kind == SGX_EOS_SPINER_RHOT) {
    set.e[k].kind = b->kind;
    set.e[k].rhot = static_cast<const SpinerRhoT *>(b->dev_struct);
  } else if (b->kind == SGX_EOS_SPINER_RHOSIE) {
    set.e[k].kind = b->kind;
    set.e[k].rhosie = static_cast<const SpinerRhoSie *>(b->dev_struct);
  } else if (b->kind >= 0 && b->kind <= SGX_EOS_DAVIS_PRODUCTS) {
    set.e[k] = b->analytic;
  } else {
    return false;
  }
  return true;
}
// the run-time accessor (TabEos kernels) is needed for tabulated and for modified materials
inline bool needs_tab_accessor(const sgx_eos_s *h) {
  return h->kind == SGX_EOS_MODIFIED || h->kind == SGX_EOS_SPINER_RHOT ||
         h->kind == SGX_EOS_SPINER_RHOSIE;
}
int launch_mod_eval(sgx_eos_s *e, int method, const double *in1, const double *in2, double *out,
                    int64_t n, double *lambda, int lambda_stride, cudaStream_t s);
int upload_tables(sgx_eos_s *e);   // cudaMalloc + H2D of e->tables.host
int ensure_staging(sgx_eos_s *e, size_t pinned_bytes, size_t device_bytes);
// Stream-ordered kernel scratch from the library's OWN memory pool (one per device, release
// threshold = never): the default pool hands its memory back to the driver at every
// synchronisation, so a caller that syncs between calls paid a fresh multi-hundred-MB cudaMalloc
// per call (measured: T(rho,sie) 3.5 ms -> 8.7 ms per 2^26 cells).  Free with cudaFreeAsync.
cudaError_t scratch_alloc_async(void **ptr, size_t bytes, cudaStream_t s);
int ensure_host_streams(sgx_eos_s *e);
int build_spiner_rhot(const sgx_sp5_tables *raw, int reproducible, int pmin_vapor_dome,
                      sgx_eos_s *e);
int build_spiner_rhosie(const sgx_sp5_tables *raw, int reproducible, int pmin_vapor_dome,
                        sgx_eos_s *e);
int build_helmholtz(const char *path, const sgx_helm_options *opt, sgx_eos_s *e);
int build_helmholtz_raw(const double *raw, size_t nraw, const sgx_helm_options *opt,
                        sgx_eos_s *e);
int launch_helm_fill(sgx_eos_s *e, double *rho, double *T, double *en, double *P, double *cv,
                     double *bmod, int64_t n, unsigned long output, double *lambda,
                     int stride, cudaStream_t s);
int launch_helm_eval(sgx_eos_s *e, int method, const double *in1, const double *in2,
                     double *out, int64_t n, double *lambda, int stride, cudaStream_t s);
int launch_pte_rhot(const EosSet &set, bool bench_mix, const MixParamsDev &prm, double *rho,
                    double *vfrac, double *sie, double *temp, double *press,
                    const double *sie_tot, const double *vfrac_tot, const double *Tguess,
                    uint64_t *status, double *residual, uint64_t *stats, int64_t ncell,
                    cudaStream_t s);
int launch_pte_fixedt(const EosSet &set, bool bench_mix, const MixParamsDev &prm, double *rho,
                      double *vfrac, double *sie, double *temp, double *press,
                      const double *sie_tot, const double *vfrac_tot, const double *Tguess,
                      uint64_t *status, double *residual, uint64_t *stats, int64_t ncell,
                      cudaStream_t s);
MixParamsDev mix_params_to_dev(const sgx_mix_params *p); // null = library defaults
// mixtures with tabulated materials (sgx_pte_tab.cu); lambda = [ncell][nmat][3] caches or null
int launch_pte_tab(int solver, const EosSet &set, const MixParamsDev &prm, double *rho,
                   double *vfrac, double *sie, double *temp, double *press,
                   const double *sie_tot, const double *vfrac_tot, const double *Tguess,
                   uint64_t *status, double *residual, uint64_t *stats, int64_t ncell,
                   cudaStream_t s, double *lambda);
int launch_pte_fixedp(const EosSet &set, bool bench_mix, const MixParamsDev &prm, double *rho,
                      double *vfrac, double *sie, double *temp, double *press,
                      const double *sie_tot, const double *vfrac_tot, const double *Tguess,
                      uint64_t *status, double *residual, uint64_t *stats, int64_t ncell,
                      cudaStream_t s);
int launch_pte_rhou(const EosSet &set, bool bench_mix, const MixParamsDev &prm, double *rho,
                    double *vfrac, double *sie, double *temp, double *press,
                    const double *sie_tot, const double *vfrac_tot, const double *Tguess,
                    uint64_t *status, double *residual, uint64_t *stats, int64_t ncell,
                    cudaStream_t s);
// Search hint of the direct T(rho, sie) inversion (sgx_spiner_root.cu).  Per density row j:
// base = sie(j, 0), s = sie(j, 1) - sie(j, 0); per pair (j, j+1), stored with row j: umin, inv_du.
// A cell's energy is binned in the coordinate of ITS blended row,
//     u = log10_nqt(max(e - base_b, 0) + s_b),  base_b = w0 base_j + w1 base_j+1,  s_b likewise,
// i.e. relative to where the blended row starts: rows of neighbouring densities differ mostly by
// a shift (the cold curve), so in this coordinate the bracketing table cell barely depends on
// the blend weight and "sie#hint"[pair][bin] = lo | hi << 16 is 3-4 cells wide.
// inv_du == 0 marks a pair that is not strictly monotone in T (no direct inversion).
struct RootRowMeta {
  double base, s, umin, inv_du;
};
constexpr int kRootHintBins = 512;
// ONE definition for the host code that fills the table and the kernel that reads it (integer
// NQT log, no FMA on either side: identical bins)
SGX_HD int root_hint_bin(const double base, const double s, const double umin,
                         const double inv_du, const double e) {
  double arg = e - base;
  arg = arg < 0.0 ? 0.0 : arg;
  const double u = fastmath::log10(arg + s);
  int k = (int)((u - umin) * inv_du);
  return k < 0 ? 0 : (k >= kRootHintBins ? kRootHintBins - 1 : k);
}

// SpinerRhoT T(rho, e): direct inversion + reference iterate for the listed cells
// (sgx_spiner_root.cu).  root_mode(): 0 hybrid (default), 1 exact, 2 direct (SGX_ROOT_MODE).
int root_mode();
int launch_rhot_T_from_rho_e_hybrid(sgx_eos_s *e, bool bounded, const double *rho,
                                    const double *sie, double *out, int64_t n, double *lambda,
                                    int ls, cudaStream_t s);
// launch geometry: persistent-ish grid-stride, multiple of the SM count
struct LaunchCfg {
  int blocks, threads;
};
LaunchCfg stream_cfg(int64_t work_items, int threads, int blocks_per_sm);
int sm_count();
} // namespace sgx

#endif // SGX_INTERNAL_HPP_
