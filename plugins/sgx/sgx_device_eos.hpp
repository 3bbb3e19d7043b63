// Reference-side binding of libsgx_b200: a singularity-eos PLUGIN (cmake: -DSINGULARITY_PLUGINS=
// plugins/sgx, cmake/plugins.cmake:20-64; precedent example/plugin/dust/dust.hpp) that joins
// singularity::Variant and routes the VECTOR overloads -- the place where the reference's hot path
// crosses from host code to the device (eos/eos_base.hpp:169-230, 503-531) -- to the B200
// library through its C ABI (include/sgx_b200.h).  Everything else (scalar calls, metadata,
// modifiers, serialization) keeps running the reference's own class, held by value as `host_`.
//
//   singularity::SgxDevice<HostEos>      HostEos = IdealGas, Gruneisen, DavisReactants,
//                                        DavisProducts, SpinerEOSDependsRhoT,
//                                        SpinerEOSDependsRhoSie, Helmholtz
//
// The override pattern is EOSPAC's (eos/eos_eospac.hpp:390-461: whole arrays handed to an external
// library, scratch ignored).  Compiled and run by tests/cpp/test_plugin.cpp against the reference
// headers where they lie, through singularity::Variant's Space-first overloads with a
// LambdaIndexer (eos/eos_variant.hpp:56-143).
//
// Memory spaces: with `PortsOfCall::Exec::Device` the arrays are DEVICE pointers on the handle's
// GPU (what Kokkos-CUDA requires of the reference's vector calls); with `Exec::Host` they are host
// arrays and the library stages them (sgx_*_host entry points: chunked H2D / kernel / D2H).
// Calls with Exec::Device are asynchronous on the CUDA stream set with set_stream() (default 0);
// callers fence as with Kokkos (test/test_eos_vector.cpp:200-203).
#ifndef SGX_PLUGIN_DEVICE_EOS_HPP_
#define SGX_PLUGIN_DEVICE_EOS_HPP_

#include <cstddef>
#include <string>
#include <type_traits>
#include <utility>

#include <ports-of-call/portability.hpp>
#include <ports-of-call/portable_errors.hpp>
#include <singularity-eos/base/constants.hpp>
#include <singularity-eos/base/variadic_utils.hpp>
#include <singularity-eos/eos/eos_base.hpp>

#include <sgx_b200.h>

namespace singularity {

namespace sgx_plugin {

// ---- LambdaIndexer -> (pointer, stride) ------------------------------------------------------
// The library addresses lambda_i = data + stride * i (the reference's own lambdaIndexer2D,
// eos/singularity_eos.cpp:38-49).  Any indexer whose operator[](i) yields a Real* that advances
// by a constant stride qualifies; the layout is probed at both ends and rejected otherwise.
struct LambdaView {
  Real *ptr;
  int stride;
};
template <typename T, typename = void>
struct yields_real_pointer : std::false_type {};
template <typename T>
struct yields_real_pointer<
    T, std::enable_if_t<std::is_convertible_v<decltype(std::declval<T &>()[0]), Real *>>>
    : std::true_type {};

template <typename LambdaIndexer>
inline LambdaView lambda_pointer_and_stride(LambdaIndexer &&lambdas, const int num,
                                            const int nlambda) {
  using L = std::remove_cv_t<std::remove_reference_t<LambdaIndexer>>;
  if constexpr (std::is_null_pointer_v<L>) {
    return LambdaView{nullptr, 0};
  } else {
    static_assert(yields_real_pointer<L>::value,
                  "sgx plugin: the lambda indexer must return Real* rows (data + stride * i); "
                  "copy type-indexed lambdas into such an array first");
    if (num <= 0 || nlambda <= 0) return LambdaView{nullptr, 0};
    Real *first = lambdas[0];
    if (first == nullptr) return LambdaView{nullptr, 0}; // NullIndexer (eos_variant.hpp:47-52)
    if (num == 1) return LambdaView{first, nlambda};
    Real *second = lambdas[1];
    const std::ptrdiff_t stride = second - first;
    Real *last = lambdas[num - 1];
    if (stride < nlambda || last != first + stride * static_cast<std::ptrdiff_t>(num - 1))
      PORTABLE_ALWAYS_THROW_OR_ABORT(
          "sgx plugin: lambda rows are not laid out as data + stride * i with stride >= nlambda");
    return LambdaView{first, static_cast<int>(stride)};
  }
}

template <typename T>
inline constexpr bool is_real_pointer_v =
    std::is_pointer_v<std::remove_reference_t<T>> &&
    std::is_same_v<std::remove_cv_t<std::remove_pointer_t<std::remove_reference_t<T>>>, Real>;

inline void check(const int rc) {
  if (rc != SGX_OK) PORTABLE_ALWAYS_THROW_OR_ABORT(sgx_last_error());
}
} // namespace sgx_plugin

using namespace eos_base;

template <typename HostEos>
class SgxDevice : public EosBase<SgxDevice<HostEos>> {
 public:
  SgxDevice() = default;
  // `dev` is a handle made by one of the sgx_eos_create_* calls from the same parameters /
  // tables as `host` (helpers in sgx_make.hpp); the object stays trivially copyable, like every
  // EOS the Variant holds, and Finalize() releases the handle.
  SgxDevice(const HostEos &host, sgx_eos_t dev) : host_(host), dev_(dev) {}

  // Tables are resident since creation: the "device" object is a second holder of the same
  // handle, finalized on its own like the reference's (sgx_eos_retain / sgx_eos_destroy).
  SgxDevice GetOnDevice() {
    if (dev_) sgx_plugin::check(sgx_eos_retain(dev_));
    return *this;
  }
  inline void Finalize() {
    if (dev_) sgx_eos_destroy(dev_);
    dev_ = nullptr;
    host_.Finalize();
  }
  void set_stream(void *cuda_stream) { stream_ = cuda_stream; }
  sgx_eos_t handle() const { return dev_; }
  const HostEos &host() const { return host_; }

  // ---- scalar methods: the reference class, unchanged ---------------------------------------
#define SGX_PLUGIN_SCALAR(NAME)                                                          \
  template <typename Indexer_t = Real *>                                                   \
  PORTABLE_INLINE_FUNCTION Real NAME(const Real a, const Real b,                           \
                                     Indexer_t &&lambda = static_cast<Real *>(nullptr))   \
      const {                                                                              \
    return host_.NAME(a, b, std::forward<Indexer_t>(lambda));                              \
  }
  SGX_PLUGIN_SCALAR(TemperatureFromDensityInternalEnergy)
  SGX_PLUGIN_SCALAR(InternalEnergyFromDensityTemperature)
  SGX_PLUGIN_SCALAR(PressureFromDensityTemperature)
  SGX_PLUGIN_SCALAR(PressureFromDensityInternalEnergy)
  SGX_PLUGIN_SCALAR(EntropyFromDensityTemperature)
  SGX_PLUGIN_SCALAR(EntropyFromDensityInternalEnergy)
  SGX_PLUGIN_SCALAR(SpecificHeatFromDensityTemperature)
  SGX_PLUGIN_SCALAR(SpecificHeatFromDensityInternalEnergy)
  SGX_PLUGIN_SCALAR(BulkModulusFromDensityTemperature)
  SGX_PLUGIN_SCALAR(BulkModulusFromDensityInternalEnergy)
  SGX_PLUGIN_SCALAR(GruneisenParamFromDensityTemperature)
  SGX_PLUGIN_SCALAR(GruneisenParamFromDensityInternalEnergy)
#undef SGX_PLUGIN_SCALAR
  template <typename Indexer_t = Real *>
  PORTABLE_INLINE_FUNCTION Real MinInternalEnergyFromDensity(
      const Real rho, Indexer_t &&lambda = static_cast<Real *>(nullptr)) const {
    return host_.MinInternalEnergyFromDensity(rho, std::forward<Indexer_t>(lambda));
  }
  template <typename Indexer_t = Real *>
  PORTABLE_INLINE_FUNCTION void
  FillEos(Real &rho, Real &temp, Real &energy, Real &press, Real &cv, Real &bmod,
          const unsigned long output, Indexer_t &&lambda = static_cast<Real *>(nullptr)) const {
    host_.FillEos(rho, temp, energy, press, cv, bmod, output, std::forward<Indexer_t>(lambda));
  }
  template <typename Indexer_t = Real *>
  PORTABLE_INLINE_FUNCTION void
  ValuesAtReferenceState(Real &rho, Real &temp, Real &sie, Real &press, Real &cv, Real &bmod,
                         Real &dpde, Real &dvdt,
                         Indexer_t &&lambda = static_cast<Real *>(nullptr)) const {
    host_.ValuesAtReferenceState(rho, temp, sie, press, cv, bmod, dpde, dvdt,
                                 std::forward<Indexer_t>(lambda));
  }
  template <typename Indexer_t = Real *>
  PORTABLE_INLINE_FUNCTION void
  DensityEnergyFromPressureTemperature(const Real press, const Real temp, Indexer_t &&lambda,
                                       Real &rho, Real &sie) const {
    host_.DensityEnergyFromPressureTemperature(press, temp, std::forward<Indexer_t>(lambda), rho,
                                               sie);
  }

  SG_ADD_BASE_CLASS_USINGS(SgxDevice<HostEos>)

  // ---- vector overloads: whole arrays to libsgx_b200 ------------------------------------------
  // Same template heads as SG_EOS_VEC_2IN_1OUT (eos_base.hpp:169-230), so these REPLACE the base
  // class's per-element loops for this type.  Raw Real pointers go to the library; any other
  // indexer type keeps the reference's loop over the (host) scalar method.
#define SGX_PLUGIN_VEC(NAME, METHOD)                                                     \
  template <typename Space, typename RealIndexer, typename ConstRealIndexer,               \
            typename LambdaIndexer,                                                        \
            typename EnableIfSpace =                                                       \
                std::enable_if_t<!variadic_utils::has_int_index_v<Space>>>                 \
  inline void NAME(const Space &s, ConstRealIndexer &&in1, ConstRealIndexer &&in2,         \
                   RealIndexer &&out, const int num, LambdaIndexer &&lambdas) const {      \
    if constexpr (sgx_plugin::is_real_pointer_v<ConstRealIndexer> &&                       \
                  sgx_plugin::is_real_pointer_v<RealIndexer>) {                            \
      vector_call<Space>(METHOD, in1, in2, out, num, std::forward<LambdaIndexer>(lambdas)); \
    } else {                                                                               \
      const SgxDevice &copy = *this;                                                       \
      portableFor(                                                                         \
          #NAME, s, 0, num,                                                                \
          PORTABLE_LAMBDA(const int i) { out[i] = copy.NAME(in1[i], in2[i], lambdas[i]); }); \
    }                                                                                      \
  }                                                                                        \
  template <typename Space, typename RealIndexer, typename ConstRealIndexer,               \
            typename LambdaIndexer,                                                        \
            typename EnableIfSpace =                                                       \
                std::enable_if_t<!variadic_utils::has_int_index_v<Space>>>                 \
  inline void NAME(const Space &s, ConstRealIndexer &&in1, ConstRealIndexer &&in2,         \
                   RealIndexer &&out, Real * /*scratch*/, const int num,                   \
                   LambdaIndexer &&lambdas, Transform && = Transform()) const {            \
    NAME(s, std::forward<ConstRealIndexer>(in1), std::forward<ConstRealIndexer>(in2),      \
         std::forward<RealIndexer>(out), num, std::forward<LambdaIndexer>(lambdas));       \
  }                                                                                        \
  template <typename RealIndexer, typename ConstRealIndexer, typename LambdaIndexer,       \
            typename EnableIfIndexed =                                                     \
                std::enable_if_t<variadic_utils::has_int_index_v<ConstRealIndexer>>>       \
  inline void NAME(ConstRealIndexer &&in1, ConstRealIndexer &&in2, RealIndexer &&out,      \
                   const int num, LambdaIndexer &&lambdas) const {                         \
    NAME(PortsOfCall::Exec::Device(), std::forward<ConstRealIndexer>(in1),                 \
         std::forward<ConstRealIndexer>(in2), std::forward<RealIndexer>(out), num,         \
         std::forward<LambdaIndexer>(lambdas));                                            \
  }                                                                                        \
  template <typename RealIndexer, typename ConstRealIndexer, typename LambdaIndexer,       \
            typename EnableIfIndexed =                                                     \
                std::enable_if_t<variadic_utils::has_int_index_v<ConstRealIndexer>>>       \
  inline void NAME(ConstRealIndexer &&in1, ConstRealIndexer &&in2, RealIndexer &&out,      \
                   Real * /*scratch*/, const int num, LambdaIndexer &&lambdas,             \
                   Transform && = Transform()) const {                                     \
    NAME(PortsOfCall::Exec::Device(), std::forward<ConstRealIndexer>(in1),                 \
         std::forward<ConstRealIndexer>(in2), std::forward<RealIndexer>(out), num,         \
         std::forward<LambdaIndexer>(lambdas));                                            \
  }
  SGX_PLUGIN_VEC(PressureFromDensityTemperature, SGX_P_FROM_RHO_T)
  SGX_PLUGIN_VEC(InternalEnergyFromDensityTemperature, SGX_E_FROM_RHO_T)
  SGX_PLUGIN_VEC(TemperatureFromDensityInternalEnergy, SGX_T_FROM_RHO_E)
  SGX_PLUGIN_VEC(PressureFromDensityInternalEnergy, SGX_P_FROM_RHO_E)
  SGX_PLUGIN_VEC(SpecificHeatFromDensityTemperature, SGX_CV_FROM_RHO_T)
  SGX_PLUGIN_VEC(SpecificHeatFromDensityInternalEnergy, SGX_CV_FROM_RHO_E)
  SGX_PLUGIN_VEC(BulkModulusFromDensityTemperature, SGX_BMOD_FROM_RHO_T)
  SGX_PLUGIN_VEC(BulkModulusFromDensityInternalEnergy, SGX_BMOD_FROM_RHO_E)
  SGX_PLUGIN_VEC(GruneisenParamFromDensityTemperature, SGX_GAMMA_FROM_RHO_T)
  SGX_PLUGIN_VEC(GruneisenParamFromDensityInternalEnergy, SGX_GAMMA_FROM_RHO_E)
#undef SGX_PLUGIN_VEC

  // FillEos vector overload (eos_base.hpp:503-531)
  template <typename Space, typename RealIndexer, typename LambdaIndexer,
            typename EnableIfSpace =
                std::enable_if_t<!variadic_utils::has_int_index_v<Space>>>
  inline void FillEos(const Space &s, RealIndexer &&rhos, RealIndexer &&temps,
                      RealIndexer &&energies, RealIndexer &&presses, RealIndexer &&cvs,
                      RealIndexer &&bmods, const int num, const unsigned long output,
                      LambdaIndexer &&lambdas) const {
    if constexpr (sgx_plugin::is_real_pointer_v<RealIndexer>) {
      const sgx_plugin::LambdaView lv =
          sgx_plugin::lambda_pointer_and_stride(std::forward<LambdaIndexer>(lambdas), num,
                                                host_.nlambda());
      if constexpr (std::is_same_v<Space, PortsOfCall::Exec::Host>) {
        sgx_plugin::check(sgx_FillEos_host(dev_, rhos, temps, energies, presses, cvs, bmods, num,
                                           output, lv.ptr, lv.stride));
      } else {
        sgx_plugin::check(sgx_FillEos(dev_, rhos, temps, energies, presses, cvs, bmods, num,
                                      output, lv.ptr, lv.stride, stream_));
      }
    } else {
      const SgxDevice &copy = *this;
      portableFor(
          "FillEos", s, 0, num, PORTABLE_LAMBDA(const int i) {
            copy.FillEos(rhos[i], temps[i], energies[i], presses[i], cvs[i], bmods[i], output,
                         lambdas[i]);
          });
    }
  }
  template <typename RealIndexer, typename LambdaIndexer,
            typename EnableIfIndexed =
                std::enable_if_t<variadic_utils::has_int_index_v<RealIndexer>>>
  inline void FillEos(RealIndexer &&rhos, RealIndexer &&temps, RealIndexer &&energies,
                      RealIndexer &&presses, RealIndexer &&cvs, RealIndexer &&bmods,
                      const int num, const unsigned long output, LambdaIndexer &&lambdas) const {
    FillEos(PortsOfCall::Exec::Device(), std::forward<RealIndexer>(rhos),
            std::forward<RealIndexer>(temps), std::forward<RealIndexer>(energies),
            std::forward<RealIndexer>(presses), std::forward<RealIndexer>(cvs),
            std::forward<RealIndexer>(bmods), num, output, std::forward<LambdaIndexer>(lambdas));
  }

  // ---- metadata -------------------------------------------------------------------------------
  PORTABLE_INLINE_FUNCTION int nlambda() const noexcept { return host_.nlambda(); }
  static constexpr unsigned long PreferredInput() { return HostEos::PreferredInput(); }
  // the library keeps its solver / lookup scratch on chip (eos_base.hpp:559-563 semantics: 0)
  static inline unsigned long scratch_size(const std::string, unsigned int) { return 0; }
  static inline unsigned long max_scratch_size(unsigned int) { return 0; }
  PORTABLE_INLINE_FUNCTION void PrintParams() const { host_.PrintParams(); }
  PORTABLE_FORCEINLINE_FUNCTION Real MinimumDensity() const { return host_.MinimumDensity(); }
  PORTABLE_FORCEINLINE_FUNCTION Real MinimumTemperature() const {
    return host_.MinimumTemperature();
  }
  PORTABLE_FORCEINLINE_FUNCTION Real MaximumDensity() const { return host_.MaximumDensity(); }
  PORTABLE_FORCEINLINE_FUNCTION Real MinimumPressure() const { return host_.MinimumPressure(); }
  PORTABLE_FORCEINLINE_FUNCTION Real MaximumPressureAtTemperature(const Real T) const {
    return host_.MaximumPressureAtTemperature(T);
  }
  PORTABLE_INLINE_FUNCTION Real MeanAtomicMass() const { return host_.MeanAtomicMass(); }
  PORTABLE_INLINE_FUNCTION Real MeanAtomicNumber() const { return host_.MeanAtomicNumber(); }
  static std::string EosType() { return std::string("Sgx") + HostEos::EosType(); }
  static std::string EosPyType() { return EosType(); }

 private:
  template <typename Space, typename LambdaIndexer>
  inline void vector_call(const int method, const Real *in1, const Real *in2, Real *out,
                          const int num, LambdaIndexer &&lambdas) const {
    const sgx_plugin::LambdaView lv = sgx_plugin::lambda_pointer_and_stride(
        std::forward<LambdaIndexer>(lambdas), num, host_.nlambda());
    if constexpr (std::is_same_v<Space, PortsOfCall::Exec::Host>) {
      sgx_plugin::check(sgx_eos_eval_host(dev_, method, in1, in2, out, num, lv.ptr, lv.stride));
    } else {
      sgx_plugin::check(
          sgx_eos_eval(dev_, method, in1, in2, out, num, lv.ptr, lv.stride, stream_));
    }
  }

  HostEos host_;
  sgx_eos_t dev_ = nullptr;
  void *stream_ = nullptr;
};

} // namespace singularity

#endif // SGX_PLUGIN_DEVICE_EOS_HPP_
