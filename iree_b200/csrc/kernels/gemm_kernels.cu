// gemm_kernels.cu — extern "C" entry points of the tcgen05 GEMM family.
//
// ABI (reference: samples/custom_dispatch/cuda/kernels/kernels.cu:7-34):
//   #hal.pipeline.layout<constants = 5, bindings = [
//     #hal.pipeline.binding<storage_buffer, ReadOnly>,   // lhs  [batch, M, K]
//     #hal.pipeline.binding<storage_buffer, ReadOnly>,   // rhs  [batch, K, N] (or [batch, N, K] for *_nt)
//     #hal.pipeline.binding<storage_buffer>]>            // out  [batch, M, N]
//   constants: M, N, K, batch, flags (bit 0: accumulate into `out`)
//   workgroup_size = [256, 1, 1]; workgroup count = persistent (see ops.c count region)
// followed by the three CUtensorMap parameters described by `<name>__tma`.

#include "gemm_sm100.cuh"
#include "iree_b200_tma_recipe.h"

using namespace b200;

#define B200_GEMM_KERNEL(NAME, CLUSTER_ATTR, ...)                                              \
  using NAME##_cfg = GemmConfig<__VA_ARGS__>;                                                  \
  extern "C" __device__ const iree_b200_tma_recipe_t NAME##__tma =                             \
      make_gemm_recipe<NAME##_cfg>();                                                          \
  extern "C" __device__ const uint32_t NAME##__smem = NAME##_cfg::kSmemBytes;                  \
  extern "C" __global__ void CLUSTER_ATTR __launch_bounds__(256, 1)                            \
      NAME(const void* __restrict__ lhs, const void* __restrict__ rhs, void* __restrict__ out, \
           int M, int N, int K, int batch, int flags,                                          \
           const __grid_constant__ TensorMap tm_a, const __grid_constant__ TensorMap tm_b,     \
           const __grid_constant__ TensorMap tm_c) {                                           \
    gemm_sm100_body<NAME##_cfg>(lhs, rhs, out, M, N, K, batch, flags, &tm_a, &tm_b, &tm_c);    \
  }

#define CLUSTER2 __cluster_dims__(2, 1, 1)
#define NOCLUSTER

// cta_group::2 (256x256 tile per SM pair), 6-stage ring.
B200_GEMM_KERNEL(iree_b200_matmul_bf16_f32, CLUSTER2, 2, KIND_F16, FMT_BF16, false, 6)
B200_GEMM_KERNEL(iree_b200_matmul_f16_f32, CLUSTER2, 2, KIND_F16, FMT_F16, false, 6)
B200_GEMM_KERNEL(iree_b200_matmul_i8_i32, CLUSTER2, 2, KIND_I8, FMT_S8, false, 6)
// transpose_rhs forms (rhs is [N, K]).
B200_GEMM_KERNEL(iree_b200_matmul_nt_bf16_f32, CLUSTER2, 2, KIND_F16, FMT_BF16, true, 6)
B200_GEMM_KERNEL(iree_b200_matmul_nt_f16_f32, CLUSTER2, 2, KIND_F16, FMT_F16, true, 6)
B200_GEMM_KERNEL(iree_b200_matmul_nt_i8_i32, CLUSTER2, 2, KIND_I8, FMT_S8, true, 6)
// cta_group::1 (128x256 tile per SM), 4-stage ring: no cluster requirement on the grid.
B200_GEMM_KERNEL(iree_b200_matmul_bf16_f32_1cta, NOCLUSTER, 1, KIND_F16, FMT_BF16, false, 4)
B200_GEMM_KERNEL(iree_b200_matmul_f16_f32_1cta, NOCLUSTER, 1, KIND_F16, FMT_F16, false, 4)
B200_GEMM_KERNEL(iree_b200_matmul_i8_i32_1cta, NOCLUSTER, 1, KIND_I8, FMT_S8, false, 4)
B200_GEMM_KERNEL(iree_b200_matmul_nt_bf16_f32_1cta, NOCLUSTER, 1, KIND_F16, FMT_BF16, true, 4)
