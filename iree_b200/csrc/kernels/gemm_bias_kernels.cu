// gemm_bias_kernels.cu — extern "C" entry points of the tcgen05 GEMMs with a fused
// bias / activation epilogue (same body as gemm_kernels.cu; GemmConfig::kBias).
//
// ABI (reference: samples/custom_dispatch/cuda/kernels/kernels.cu:7-34):
//   #hal.pipeline.layout<constants = 5, bindings = [
//     #hal.pipeline.binding<storage_buffer, ReadOnly>,   // lhs  [batch, M, K]
//     #hal.pipeline.binding<storage_buffer, ReadOnly>,   // rhs  [batch, K, N] (or [batch, N, K] for *_nt)
//     #hal.pipeline.binding<storage_buffer, ReadOnly>,   // bias [N], accumulator type (f32 / i32)
//     #hal.pipeline.binding<storage_buffer>]>            // out  [batch, M, N]
//   constants: M, N, K, batch, flags (bit 0: accumulate into `out` first; bit 1: then
//   max(x, 0) after the bias — arith.maximumf / arith.maxsi)
//   out[b, m, n] = act((out[b, m, n] +) sum_k lhs·rhs + bias[n]), optionally truncated to
//   the operand type (round to nearest even) by the *_bf16_bf16 / *_f16_f16 exports.
// followed by the three CUtensorMap parameters described by `<name>__tma`.

#include "gemm_sm100.cuh"
#include "iree_b200_tma_recipe.h"

using namespace b200;

#define B200_CLUSTER_ATTR_1
#define B200_CLUSTER_ATTR_2 __cluster_dims__(2, 1, 1)

#define X(NAME, CLUSTER, ...)                                                                  \
  using NAME##_cfg = GemmConfig<__VA_ARGS__>;                                                  \
  extern "C" __device__ const iree_b200_tma_recipe_t NAME##__tma =                             \
      make_gemm_recipe<NAME##_cfg>();                                                          \
  extern "C" __global__ void B200_CLUSTER_ATTR_##CLUSTER __launch_bounds__(256, 1)             \
      NAME(const void* __restrict__ lhs, const void* __restrict__ rhs,                         \
           const void* __restrict__ bias, void* __restrict__ out, int M, int N, int K,         \
           int batch, int flags, const __grid_constant__ TensorMap tm_a,                       \
           const __grid_constant__ TensorMap tm_b, const __grid_constant__ TensorMap tm_c) {   \
    gemm_sm100_body<NAME##_cfg>(lhs, rhs, out, M, N, K, batch, flags, &tm_a, &tm_b, &tm_c,     \
                                bias);                                                         \
  }
#include "gemm_bias_exports.inc"
#undef X
