// mmt4d_kernels.cu — linalg.mmt4d on data-tiled operands, SIMT form.
//
// PENDING VALIDATION: written after this round's GPU budget was spent. The device code below
// has run only on host threads (tests/mock_cuda/mock_kernels_mmt4d.cc under the mock driver:
// bit-exact against the oracle and against the reference's own iree_uk_mmt4d, also under
// ASAN). Ships in the on-demand second executable (build.py EXPERIMENTAL_UNITS).
//
// IREE's encoding path pre-packs matmul operands into tiles ("data tiling"); the contraction
// on packed operands is linalg.mmt4d, whose CPU lowering is the iree_uk_mmt4d micro-kernel
// (runtime/src/iree/builtins/ukernel/mmt4d.h:14-22 parameters, mmt4d.c:55-103 outer loops,
// mmt4d_tile_generic.c:54-79,199-227,289-317 tiles; element-wise statement
// tools/mmt4d_test.c:17-31):
//   lhs [M1, K1, M0, K0], rhs [N1, K1, N0, K0] (the "t": rhs tiles are N-major),
//   out [M1, N1, M0, N0];
//   out[m1,n1,m0,n0] = (accumulate ? out[m1,n1,m0,n0] : 0)
//                      + sum over k1 ascending, k0 ascending of lhs[m1,k1,m0,k0] * rhs[n1,k1,n0,k0]
// in f32 (f16 / bf16 operands) or wrapping i32 (i8 operands). Tile sizes are run-time values:
// the compiler picks them per target (16x16x2 for AVX-512-BF16, 8x8x1, 16x16x16 for a GPU MMA
// intrinsic, …), so a module packed for any of them runs unchanged.
//
// ABI: bindings lhs, rhs (read-only), out; constants M1, N1, K1, M0, N0, K0, flags (bit 0:
// accumulate into out). workgroup_size = [256,1,1]; workgroup count =
// [ceil(N1*N0/64), ceil(M1*M0/64), 1]: a block owns a 64x64 region of the logical
// [M1*M0, N1*N0] result, k accumulated in ascending order exactly as the reference loops do,
// so THIS kernel's results are bit-identical to iree_uk_mmt4d's generic tiles for any input
// whose f32 products are exact (all f16 / bf16 products are) and for every i8 input.
// The tensor-core route the ops layer takes for tile-filling shapes (relayout -> tcgen05 GEMM
// -> relayout, ops.cc) accumulates in f32 in the tensor core's order instead: bit-identical
// for i8 and for inputs whose partial sums are exact (the reference generator's), within
// f32 summation-order error (<= 1e-5 of the result's scale; tests/test_zz_mmt4d_gpu.py::
// test_tensor_route_vs_simt_on_random_data) otherwise. IREE_B200_MATMUL_FLAG_FORCE_GENERIC
// keeps a call on this kernel.

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>

// ---- tcgen05 form: tiles read and written directly -----------------------------------------
// iree_b200_mmt4d_{bf16,f16}_f32 / _i8_i32 [_1cta]: for tiles whose K0 row is 16 bytes (K0 = 8
// for f16 / bf16, 16 for i8) and M0, N0 in {8, 16, 32, 64, 128}, the operand tiles reach shared
// memory as UMMA core matrices through rank-4 tensor maps (gemm_sm100.cuh, kTiled) and the
// result goes back as tiles the same way: ONE dispatch, no workspace, the tcgen05 GEMM's rate.
// Bindings lhs, rhs (read-only), out; constants M1, N1, K1, M0, N0, K0, flags (bit 0
// accumulate), then the map extents make_gemm_tiled_recipe lists (M0 / 8, 128 / M0, N0 / 8,
// (256 or 128) / N0, min(N0, 32), 32 / min(N0, 32), result swizzle mode); then the three
// tensor maps. Device-only (the SIMT kernels below also compile for the host shims).
#if defined(__CUDACC__)
#include "gemm_sm100.cuh"

#define B200_MMT4D(NAME, CLUSTER, G, KIND, FMT, STAGES)                                          \
  using NAME##_cfg = b200::GemmConfig<G, b200::KIND, b200::FMT, true, STAGES, false, false, true>; \
  extern "C" __device__ const iree_b200_tma_recipe_t NAME##__tma =                              \
      b200::make_gemm_tiled_recipe<NAME##_cfg>();                                               \
  extern "C" __global__ void CLUSTER __launch_bounds__(256, 1)                                  \
      NAME(const void* __restrict__ lhs, const void* __restrict__ rhs, void* __restrict__ out,  \
           int M1, int N1, int K1, int M0, int N0, int K0, int flags, int m0_groups,            \
           int box_m1, int n0_groups, int box_n1, int box_c0, int box_c1, int c_swizzle,        \
           const __grid_constant__ b200::TensorMap tm_a,                                        \
           const __grid_constant__ b200::TensorMap tm_b,                                        \
           const __grid_constant__ b200::TensorMap tm_c) {                                      \
    (void)m0_groups, (void)box_m1, (void)n0_groups, (void)box_n1, (void)box_c0, (void)box_c1;   \
    b200::gemm_sm100_body<NAME##_cfg>(lhs, rhs, out, M1 * M0, N1 * N0, K1 * K0, 1, flags & 0xf1, \
                                      &tm_a, &tm_b, &tm_c, nullptr, M0, N0,                     \
                                      (1u << c_swizzle) - 1u);                                  \
  }

B200_MMT4D(iree_b200_mmt4d_bf16_f32, __cluster_dims__(2, 1, 1), 2, KIND_F16, FMT_BF16, 6)
B200_MMT4D(iree_b200_mmt4d_f16_f32, __cluster_dims__(2, 1, 1), 2, KIND_F16, FMT_F16, 6)
B200_MMT4D(iree_b200_mmt4d_i8_i32, __cluster_dims__(2, 1, 1), 2, KIND_I8, FMT_S8, 6)
B200_MMT4D(iree_b200_mmt4d_bf16_f32_1cta, , 1, KIND_F16, FMT_BF16, 4)
B200_MMT4D(iree_b200_mmt4d_f16_f32_1cta, , 1, KIND_F16, FMT_F16, 4)
B200_MMT4D(iree_b200_mmt4d_i8_i32_1cta, , 1, KIND_I8, FMT_S8, 4)
#endif  // __CUDACC__

namespace {

constexpr int kMmt4dTile = 64;
constexpr int kMmt4dTileK = 16;

__device__ __forceinline__ float mmt4d_to_acc(__nv_bfloat16 v) { return __bfloat162float(v); }
__device__ __forceinline__ float mmt4d_to_acc(__half v) { return __half2float(v); }
__device__ __forceinline__ int32_t mmt4d_to_acc(int8_t v) { return (int32_t)v; }
__device__ __forceinline__ float mmt4d_to_acc(float v) { return v; }

// Element (r, k) of a packed [R1, K1, R0, K0] operand.
__device__ __forceinline__ size_t packed_at(int r, int k, int K1, int R0, int K0) {
  return ((size_t)(r / R0) * K1 + k / K0) * ((size_t)R0 * K0) + (size_t)(r % R0) * K0 + k % K0;
}

template <typename TIn, typename TAcc>
__device__ __forceinline__ void mmt4d_generic_body(const TIn* __restrict__ lhs,
                                                   const TIn* __restrict__ rhs,
                                                   TAcc* __restrict__ out, int M1, int N1, int K1,
                                                   int M0, int N0, int K0, int flags) {
  __shared__ TAcc a_tile[kMmt4dTileK][kMmt4dTile + 1];
  __shared__ TAcc b_tile[kMmt4dTileK][kMmt4dTile + 1];

  const bool accumulate = (flags & 1) != 0;
  const int M = M1 * M0, N = N1 * N0, K = K1 * K0;  // logical (padded) extents
  const int m0 = blockIdx.y * kMmt4dTile;
  const int n0 = blockIdx.x * kMmt4dTile;
  const int tx = threadIdx.x % 16;
  const int ty = threadIdx.x / 16;

  TAcc acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int gm = m0 + ty + 16 * i, gn = n0 + tx + 16 * j;
      acc[i][j] = (accumulate && gm < M && gn < N) ? out[packed_at(gm, gn, N1, M0, N0)] : TAcc(0);
    }
  }

  for (int k0 = 0; k0 < K; k0 += kMmt4dTileK) {
    // Consecutive threads walk k: contiguous within a tile row when K0 > 1.
    for (int idx = threadIdx.x; idx < kMmt4dTile * kMmt4dTileK; idx += 256) {
      const int kk = idx % kMmt4dTileK, rr = idx / kMmt4dTileK;
      const int gk = k0 + kk, gm = m0 + rr, gn = n0 + rr;
      a_tile[kk][rr] =
          (gm < M && gk < K) ? mmt4d_to_acc(lhs[packed_at(gm, gk, K1, M0, K0)]) : TAcc(0);
      b_tile[kk][rr] =
          (gn < N && gk < K) ? mmt4d_to_acc(rhs[packed_at(gn, gk, K1, N0, K0)]) : TAcc(0);
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < kMmt4dTileK; ++kk) {
      TAcc a[4], bb[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = a_tile[kk][ty + 16 * i];
#pragma unroll
      for (int j = 0; j < 4; ++j) bb[j] = b_tile[kk][tx + 16 * j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] += a[i] * bb[j];
    }
    __syncthreads();
  }

#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int gm = m0 + ty + 16 * i;
    if (gm >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int gn = n0 + tx + 16 * j;
      if (gn >= N) continue;
      out[packed_at(gm, gn, N1, M0, N0)] = acc[i][j];
    }
  }
}

}  // namespace

extern "C" __global__ void __launch_bounds__(256)
    iree_b200_mmt4d_generic_bf16_f32(const __nv_bfloat16* __restrict__ lhs,
                                     const __nv_bfloat16* __restrict__ rhs, float* __restrict__ out,
                                     int M1, int N1, int K1, int M0, int N0, int K0, int flags) {
  mmt4d_generic_body<__nv_bfloat16, float>(lhs, rhs, out, M1, N1, K1, M0, N0, K0, flags);
}

extern "C" __global__ void __launch_bounds__(256)
    iree_b200_mmt4d_generic_f16_f32(const __half* __restrict__ lhs, const __half* __restrict__ rhs,
                                    float* __restrict__ out, int M1, int N1, int K1, int M0, int N0,
                                    int K0, int flags) {
  mmt4d_generic_body<__half, float>(lhs, rhs, out, M1, N1, K1, M0, N0, K0, flags);
}

// f32f32f32 (mmt4d_tile_generic.c's most common data-tiled case): `acc += a * b` contracts to
// an FMA here as in the reference's AVX tiles (_mm256_fmadd_ps); the generic C tile rounds the
// product first. Identical on inputs whose products are exact; <= 1 ulp per step otherwise.
extern "C" __global__ void __launch_bounds__(256)
    iree_b200_mmt4d_generic_f32_f32(const float* __restrict__ lhs, const float* __restrict__ rhs,
                                    float* __restrict__ out, int M1, int N1, int K1, int M0, int N0,
                                    int K0, int flags) {
  mmt4d_generic_body<float, float>(lhs, rhs, out, M1, N1, K1, M0, N0, K0, flags);
}

extern "C" __global__ void __launch_bounds__(256)
    iree_b200_mmt4d_generic_i8_i32(const int8_t* __restrict__ lhs, const int8_t* __restrict__ rhs,
                                   int32_t* __restrict__ out, int M1, int N1, int K1, int M0, int N0,
                                   int K0, int flags) {
  mmt4d_generic_body<int8_t, int32_t>(lhs, rhs, out, M1, N1, K1, M0, N0, K0, flags);
}

// ---- tiled <-> row-major relayout (the dispatches around the tensor-core mmt4d route) -------
//
// The tcgen05 GEMM consumes dense row-major operands; until a kernel reads tiles directly
//  the ops layer unpacks lhs / rhs (and the accumulator) into a workspace,
// runs the validated NT GEMM, and packs the result back. Pure data movement, so exact.
//   tiled [R1, C1, R0, C0]  <->  row-major [R1*R0, C1*C0]
// HBM-bound: one read and one write per element, BOTH sides coalesced. For a fixed r1 a run of
// tiles [c1, c1 + t) is contiguous in the tiled buffer (t * R0 * C0 elements) and covers R0
// row segments of t * C0 contiguous elements in the row-major buffer. A block moves one such
// panel through shared memory: linear on the tiled side, row by row on the other (the
// round-1 kernel moved one element per thread straight through and touched the tiled side in
// runs of C0 elements: 1/8 sector efficiency for a 16x16x2 bf16 tile). Tiles too large for
// the staging buffer (R0 * C0 > kPanelElems) fall back to the element-wise path.
//
// ABI: bindings src (read-only), dst; constants R1, C1, R0, C0, to_tiled (0: tiled ->
// row-major, 1: row-major -> tiled). workgroup_size = [256,1,1]; any 1-D workgroup count
// (work items are strided over the grid; the ops layer launches min(items, 8 per SM)).
constexpr int kPanelBytes = 16384;

template <typename T>
__device__ __forceinline__ void mmt4d_relayout_body(const T* __restrict__ src, T* __restrict__ dst,
                                                    int R1, int C1, int R0, int C0, int to_tiled) {
  constexpr int kPanelElems = kPanelBytes / (int)sizeof(T);
  __shared__ T panel[kPanelElems];
  const int tile = R0 * C0;
  const size_t cols = (size_t)C1 * C0;
  if (tile > kPanelElems) {
    // Element-wise fallback (huge tiles): consecutive threads along the row-major row.
    const size_t total = (size_t)R1 * R0 * cols;
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (size_t)gridDim.x * 256) {
      const int r = (int)(i / cols), c = (int)(i % cols);
      const size_t tiled = packed_at(r, c, C1, R0, C0);
      if (to_tiled) {
        dst[tiled] = src[i];
      } else {
        dst[i] = src[tiled];
      }
    }
    return;
  }
  const int tiles_per_panel = kPanelElems / tile;                     // >= 1
  const int panels_per_row = (C1 + tiles_per_panel - 1) / tiles_per_panel;
  const long long items = (long long)R1 * panels_per_row;
  for (long long w = blockIdx.x; w < items; w += gridDim.x) {
    const int r1 = (int)(w / panels_per_row);
    const int t0 = (int)(w % panels_per_row) * tiles_per_panel;
    const int t = tiles_per_panel < C1 - t0 ? tiles_per_panel : C1 - t0;
    const int n = t * tile;               // elements of the panel
    const int seg = t * C0;               // contiguous row-major elements per row
    const size_t tiled_base = ((size_t)r1 * C1 + t0) * tile;
    const size_t flat_base = (size_t)r1 * R0 * cols + (size_t)t0 * C0;
    if (to_tiled) {
      for (int j = threadIdx.x; j < n; j += 256) {
        const int r0 = j / seg, x = j % seg;
        panel[(x / C0) * tile + r0 * C0 + (x % C0)] = src[flat_base + (size_t)r0 * cols + x];
      }
      __syncthreads();
      for (int j = threadIdx.x; j < n; j += 256) dst[tiled_base + j] = panel[j];
    } else {
      for (int j = threadIdx.x; j < n; j += 256) panel[j] = src[tiled_base + j];
      __syncthreads();
      for (int j = threadIdx.x; j < n; j += 256) {
        const int r0 = j / seg, x = j % seg;
        dst[flat_base + (size_t)r0 * cols + x] = panel[(x / C0) * tile + r0 * C0 + (x % C0)];
      }
    }
    __syncthreads();
  }
}

extern "C" __global__ void __launch_bounds__(256)
    iree_b200_mmt4d_relayout_b8(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, int R1,
                                int C1, int R0, int C0, int to_tiled) {
  mmt4d_relayout_body<uint8_t>(src, dst, R1, C1, R0, C0, to_tiled);
}

extern "C" __global__ void __launch_bounds__(256)
    iree_b200_mmt4d_relayout_b16(const uint16_t* __restrict__ src, uint16_t* __restrict__ dst,
                                 int R1, int C1, int R0, int C0, int to_tiled) {
  mmt4d_relayout_body<uint16_t>(src, dst, R1, C1, R0, C0, to_tiled);
}

extern "C" __global__ void __launch_bounds__(256)
    iree_b200_mmt4d_relayout_b32(const uint32_t* __restrict__ src, uint32_t* __restrict__ dst,
                                 int R1, int C1, int R0, int C0, int to_tiled) {
  mmt4d_relayout_body<uint32_t>(src, dst, R1, C1, R0, C0, to_tiled);
}
