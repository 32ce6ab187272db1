// gemm_dtm_kernels.cu — the tcgen05 GEMM family launchable by an UNMODIFIED IREE CUDA driver.
//
// The host-fed exports (gemm_kernels.cu) take three trailing CUtensorMap parameters that the
// launch layer encodes; IREE's stock driver passes only binding pointers and i32 push constants
// (stream_command_buffer.c:532-601). These exports have exactly that ABI —
//   bindings lhs, rhs (read-only), out; constants M, N, K, batch, flags
// — and build their own tensor maps: warp 3 of every CTA copies the three templates of its
// configuration (tensormap_templates.inc: the host encoder's output for a dummy shape) to
// shared memory, patches address / extents / strides with tensormap.replace, and publishes
// them to the CTA's slots in |iree_b200_dtm_slots| (tensormap.cp_fenceproxy, release); the
// TMA-issuing threads acquire them after the kernel's setup barrier. ~2 us per launch.
//
// The slots are module globals indexed by the SM the CTA runs on: these kernels take more
// than half an SM's shared memory, so an SM holds one of their CTAs at a time and a slot is
// never shared, whatever else is in flight. (iree_b200_ops selects these exports only under
// IREE_B200_MATMUL_FLAG_DEVICE_MAPS: with the thin launch layer the host-fed twins save the
// ~2 us.)

#include "gemm_sm100.cuh"

using namespace b200;

#define B200_CLUSTER_ATTR_1
#define B200_CLUSTER_ATTR_2 __cluster_dims__(2, 1, 1)

__device__ __align__(128) TensorMap iree_b200_dtm_slots[kTensorMapSlotSms][4];  // [sm][A, B, C, pad]

// (Variadic: the three brace-enclosed rows contain commas.)
#define B200_TENSORMAP_TEMPLATE(NAME, ...) \
  __device__ __align__(128) const uint64_t NAME##__tmt[3][16] = {__VA_ARGS__};
#include "tensormap_templates.inc"
#undef B200_TENSORMAP_TEMPLATE

template <class Cfg>
__device__ __forceinline__ void build_gemm_maps(TensorMap* slots, TensorMap* staging,
                                                const uint64_t (*tmpl)[16], const void* lhs,
                                                const void* rhs, void* out, int M, int N, int K,
                                                int batch) {
  // The extents make_gemm_recipe<Cfg>() names: A [batch, M, K]; B [batch, K, N] or
  // [batch, N, K]; C [batch, M, N] of 32-bit elements (16-bit for the truncating exports).
  tensormap_build_rank3(slots + 0, staging, tmpl[0], lhs, K, M, batch, Cfg::kElemSize);
  if constexpr (Cfg::kBMajorK) {
    tensormap_build_rank3(slots + 1, staging + 1, tmpl[1], rhs, K, N, batch, Cfg::kElemSize);
  } else {
    tensormap_build_rank3(slots + 1, staging + 1, tmpl[1], rhs, N, K, batch, Cfg::kElemSize);
  }
  tensormap_build_rank3(slots + 2, staging + 2, tmpl[2], out, N, M, batch, Cfg::kOut16 ? 2 : 4);
}

// One CTA per SM (what makes the per-SM slots safe): more than half the SM's shared memory.
// Every host-fed export of gemm_exports.inc gets a twin `<name>_dtm` (B200_DTM_BIAS: the
// fused bias / activation exports of gemm_bias_exports.inc, gemm_bias_dtm_kernels.cu).
#if !defined(B200_DTM_BIAS)
#define X(NAME, CLUSTER, ...)                                                                  \
  using NAME##_cfg = GemmConfig<__VA_ARGS__>;                                                  \
  static_assert(NAME##_cfg::kSmemBytes > 232448 / 2, "per-SM tensor-map slots need 1 CTA/SM"); \
  extern "C" __global__ void B200_CLUSTER_ATTR_##CLUSTER __launch_bounds__(256, 1)             \
      NAME##_dtm(const void* __restrict__ lhs, const void* __restrict__ rhs,                   \
                 void* __restrict__ out, int M, int N, int K, int batch, int flags) {          \
    __shared__ __align__(128) TensorMap staging[3];                                            \
    TensorMap* slots = iree_b200_dtm_slots[tensormap_slot_index()];                            \
    if ((threadIdx.x >> 5) == 3) {                                                             \
      build_gemm_maps<NAME##_cfg>(slots, staging, NAME##__tmt, lhs, rhs, out, M, N, K, batch); \
    }                                                                                          \
    gemm_sm100_body<NAME##_cfg, true>(lhs, rhs, out, M, N, K, batch, flags, slots, slots + 1,  \
                                      slots + 2);                                              \
  }
#include "gemm_exports.inc"
#undef X
#else
// bindings lhs, rhs, bias (read-only), out; constants M, N, K, batch, flags (bit 1: relu).
#define X(NAME, CLUSTER, ...)                                                                  \
  using NAME##_cfg = GemmConfig<__VA_ARGS__>;                                                  \
  static_assert(NAME##_cfg::kSmemBytes > 232448 / 2, "per-SM tensor-map slots need 1 CTA/SM"); \
  extern "C" __global__ void B200_CLUSTER_ATTR_##CLUSTER __launch_bounds__(256, 1)             \
      NAME##_dtm(const void* __restrict__ lhs, const void* __restrict__ rhs,                   \
                 const void* __restrict__ bias, void* __restrict__ out, int M, int N, int K,   \
                 int batch, int flags) {                                                       \
    __shared__ __align__(128) TensorMap staging[3];                                            \
    TensorMap* slots = iree_b200_dtm_slots[tensormap_slot_index()];                            \
    if ((threadIdx.x >> 5) == 3) {                                                             \
      build_gemm_maps<NAME##_cfg>(slots, staging, NAME##__tmt, lhs, rhs, out, M, N, K, batch); \
    }                                                                                          \
    gemm_sm100_body<NAME##_cfg, true>(lhs, rhs, out, M, N, K, batch, flags, slots, slots + 1,  \
                                      slots + 2, bias);                                        \
  }
#include "gemm_bias_exports.inc"
#undef X
#endif
