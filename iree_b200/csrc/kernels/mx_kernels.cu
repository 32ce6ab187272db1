// mx_kernels.cu — OCP microscaling (MXFP4) matmul, shape-agnostic SIMT form.
//
// STATUS: compiled for sm_100a, NOT YET RUN ON A GPU (written after this round's GPU budget
// was spent). It ships in its own executable (iree_b200_pending_kernels.cda1) that the ops layer
// loads only when an MX entry point is called, so nothing on the validated paths depends on
// it; tests/test_zz_mxfp4_gpu.py is skipped unless IREE_B200_PENDING_VALIDATION=1.
//
// Semantics (tests/e2e/matmul/generate_code_mx.py:153-172; oracle.c oracle_matmul_mxfp4_f32):
//   out[m, n] = (out[m, n] +) sum over blocks kb, elements i of
//               (f32(lhs[m, kb, i]) * 2^(lhs_scale[m, kb] - 127)) *
//               (f32(rhs[n, kb, i]) * 2^(rhs_scale[n, kb] - 127))
// lhs / rhs are f4E2M1FN packed two per byte (element 2i in the low nibble,
// mmt4d_tile_generic.c:38-45), 32 elements per block, one f8E8M0FNU scale per (row, block)
// (0xFF = NaN); both operands are [rows, K] (the transpose_rhs form, the only one the
// reference generates). Every scaled operand and every product is exact in f32, so the
// result is bit-identical to the reference for any input whose partial sums are exact, and
// within f32 summation error otherwise (same k-ascending order).
//
// ABI (samples/custom_dispatch/cuda/kernels/kernels.cu:7-34): bindings lhs [M, K/2] u8,
// lhs_scales [M, K/32] u8, rhs [N, K/2] u8, rhs_scales [N, K/32] u8, out [M, N] f32;
// constants M, N, K, batch (must be 1), flags (bit 0: accumulate into out).
// workgroup_size = [256,1,1], workgroup count = [ceil(N/64), ceil(M/64), 1].

#include <stdint.h>

// The tcgen05 forms are device-only; the SIMT kernels below (and the scale repack) are also
// compiled for the host by the mock-driver tests (tests/mock_cuda/mock_kernels_mx.cc).
#if defined(__CUDACC__)
#include "mx_gemm_sm100.cuh"

// ---- tcgen05 block-scaled form ---------------------------------------------------------------
// iree_b200_matmul_mxfp4_f32 (SM pairs, 256 x 256 tiles) / _1cta (128 x 256): bindings lhs
// [M, K/2] u8, lhs_scales [M, K/32] u8, rhs [N, K/2] u8, rhs_scales [N, K/32] u8 (read-only),
// out [M, N] f32; constants M, N, K, flags (bit 0: accumulate), K / 2; then the three tensor
// maps of the recipe (lhs bytes, rhs bytes, out). workgroup_size = [384,1,1]; workgroup count
// = [min(tiles x CTAs per tile, #SMs), 1, 1] (persistent). Guards (ops layer): K % 256 == 0,
// N % 4 == 0, 16-byte aligned operand / result bindings, 4-byte aligned scale bindings.
#define B200_MXFP4(NAME, G, CLUSTER)                                                            \
  extern "C" __device__ const iree_b200_tma_recipe_t NAME##__tma =                              \
      b200::make_mx_gemm_recipe(false);                                                         \
  extern "C" __global__ void CLUSTER __launch_bounds__(b200::MxGemmConfig<G>::kThreads, 1)      \
      NAME(const void* __restrict__ lhs, const void* __restrict__ lhs_scales,                   \
           const void* __restrict__ rhs, const void* __restrict__ rhs_scales,                   \
           void* __restrict__ out, int M, int N, int K, int flags, int k_bytes,                 \
           const __grid_constant__ b200::TensorMap tm_a,                                        \
           const __grid_constant__ b200::TensorMap tm_b,                                        \
           const __grid_constant__ b200::TensorMap tm_c) {                                      \
    (void)lhs;                                                                                  \
    (void)rhs;                                                                                  \
    (void)k_bytes;                                                                              \
    b200::mx_gemm_sm100_body<b200::MxGemmConfig<G>, false>(                                     \
        static_cast<const uint8_t*>(lhs_scales), static_cast<const uint8_t*>(rhs_scales), out,  \
        M, N, K, flags, &tm_a, &tm_b, &tm_c);                                                   \
  }

B200_MXFP4(iree_b200_matmul_mxfp4_f32, 2, __cluster_dims__(2, 1, 1))
B200_MXFP4(iree_b200_matmul_mxfp4_f32_1cta, 1, )

// iree_b200_matmul_mxfp4_packed_f32[_1cta]: the same with the scale bindings already in atom
// order ([ceil(rows / 128)][K / 128][512 bytes], 16-byte aligned); constants M, N, K, flags,
// K / 2, K / 128, ceil(M / 128), ceil(N / 128); five tensor maps (lhs, rhs, out, lhs scale
// atoms, rhs scale atoms).
#define B200_MXFP4_PACKED(NAME, G, CLUSTER)                                                     \
  extern "C" __device__ const iree_b200_tma_recipe_t NAME##__tma =                              \
      b200::make_mx_gemm_recipe(true);                                                          \
  extern "C" __global__ void CLUSTER __launch_bounds__(b200::MxGemmConfig<G>::kThreads, 1)      \
      NAME(const void* __restrict__ lhs, const void* __restrict__ lhs_scales,                   \
           const void* __restrict__ rhs, const void* __restrict__ rhs_scales,                   \
           void* __restrict__ out, int M, int N, int K, int flags, int k_bytes, int k_atoms,    \
           int m_tiles, int n_tiles, const __grid_constant__ b200::TensorMap tm_a,              \
           const __grid_constant__ b200::TensorMap tm_b,                                        \
           const __grid_constant__ b200::TensorMap tm_c,                                        \
           const __grid_constant__ b200::TensorMap tm_sfa,                                      \
           const __grid_constant__ b200::TensorMap tm_sfb) {                                    \
    (void)lhs;                                                                                  \
    (void)rhs;                                                                                  \
    (void)k_bytes;                                                                              \
    (void)k_atoms;                                                                              \
    (void)m_tiles;                                                                              \
    (void)n_tiles;                                                                              \
    b200::mx_gemm_sm100_body<b200::MxGemmConfig<G>, true>(                                      \
        static_cast<const uint8_t*>(lhs_scales), static_cast<const uint8_t*>(rhs_scales), out,  \
        M, N, K, flags, &tm_a, &tm_b, &tm_c, &tm_sfa, &tm_sfb);                                 \
  }

B200_MXFP4_PACKED(iree_b200_matmul_mxfp4_packed_f32, 2, __cluster_dims__(2, 1, 1))
B200_MXFP4_PACKED(iree_b200_matmul_mxfp4_packed_f32_1cta, 1, )
#endif  // __CUDACC__

// iree_b200_mxfp4_pack_scales: row-major scales [rows, K / 32] -> the atoms the block-scaled
// instruction reads: out[(rt * (K / 128) + ka) * 512 + (r % 32) * 16 + ((r % 128) / 32) * 4 + c]
// = in[r, 4 ka + c], r = 128 rt + ..., rows padded to a multiple of 128 with the scale 1.0
// (0x7F; padded rows only reach result rows / columns the TMA store clips). One thread per
// output word (coalesced 4-byte stores; the reads are 4 aligned bytes of one row).
// bindings: scales (read-only), atoms; constants rows, K. workgroup_size [256,1,1],
// workgroup count [ceil(ceil(rows / 128) * (K / 128) * 128 / 256), 1, 1]. K % 128 == 0.
extern "C" __global__ void __launch_bounds__(256)
    iree_b200_mxfp4_pack_scales(const uint8_t* __restrict__ scales, uint32_t* __restrict__ atoms,
                                int rows, int K) {
  const int k_atoms = K / 128;
  const long long word = (long long)blockIdx.x * 256 + threadIdx.x;
  const long long total = (long long)((rows + 127) / 128) * k_atoms * 128;
  if (word >= total) return;
  const int j = (int)(word % 128);
  const long long atom = word / 128;
  const int ka = (int)(atom % k_atoms), rt = (int)(atom / k_atoms);
  const int r = rt * 128 + (j % 4) * 32 + j / 4;
  uint32_t v = 0x7F7F7F7Fu;
  if (r < rows) {
    const uint8_t* p = scales + (size_t)r * (K / 32) + ka * 4;
    v = (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
  }
  atoms[word] = v;
}

namespace {

constexpr int kTile = 64;
constexpr int kBlock = 32;  // MX block size = k elements per step

// f4E2M1FN -> f32 (base/internal/math.h:606-613): S.EE.M, bias 1, no Inf / NaN;
// exponent 0 is subnormal (0 or 0.5).
__device__ __forceinline__ float f4e2m1fn_to_f32(uint32_t code) {
  const uint32_t sign = (code & 8u) << 28;
  const uint32_t e = (code >> 1) & 3u, m = code & 1u;
  if (e == 0) return __uint_as_float(sign | (m ? 0x3f000000u : 0u));
  return __uint_as_float(sign | ((e + 126u) << 23) | (m << 22));
}

// f8E8M0FNU -> f32 (math.h:615-624): 2^(b - 127); 0xFF is NaN; b == 0 is the f32
// subnormal 2^-127.
__device__ __forceinline__ float f8e8m0fnu_to_f32(uint32_t b) {
  if (b == 0xFFu) return __uint_as_float(0x7fc00000u);
  if (b == 0u) return __uint_as_float(0x00400000u);
  return __uint_as_float(b << 23);
}

// Loads one [64 rows, 32 elements] block of a packed operand, scales it, and stores it
// k-major into |tile|. Thread t handles row t / 4, elements (t % 4) * 8 .. + 8.
__device__ __forceinline__ void stage_block(const uint8_t* __restrict__ bytes,
                                            const uint8_t* __restrict__ scales, int rows, int K,
                                            int row0, int k0, float (*tile)[kTile + 1]) {
  const int r = threadIdx.x >> 2, part = threadIdx.x & 3;
  const int row = row0 + r;
  float scale = 0.f;
  uint32_t packed = 0;
  if (row < rows) {
    scale = f8e8m0fnu_to_f32(scales[(size_t)row * (K / kBlock) + k0 / kBlock]);
    const uint8_t* p = bytes + ((size_t)row * K + k0) / 2 + part * 4;
    // Byte loads: binding pointers are only element-aligned.
    packed = (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float v = f4e2m1fn_to_f32((packed >> (4 * i)) & 0xFu);
    // Rows past the end contribute exact zeros (scale 0), never NaN.
    tile[part * 8 + i][r] = row < rows ? v * scale : 0.f;
  }
}

}  // namespace

extern "C" __global__ void __launch_bounds__(256)
    iree_b200_matmul_generic_mxfp4_f32(const uint8_t* __restrict__ lhs,
                                       const uint8_t* __restrict__ lhs_scales,
                                       const uint8_t* __restrict__ rhs,
                                       const uint8_t* __restrict__ rhs_scales,
                                       float* __restrict__ out, int M, int N, int K, int batch,
                                       int flags) {
  (void)batch;
  __shared__ float a_tile[kBlock][kTile + 1];
  __shared__ float b_tile[kBlock][kTile + 1];
  const bool accumulate = (flags & 1) != 0;
  const int m0 = blockIdx.y * kTile, n0 = blockIdx.x * kTile;
  const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;

  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int gm = m0 + ty + 16 * i, gn = n0 + tx + 16 * j;
      acc[i][j] = (accumulate && gm < M && gn < N) ? out[(size_t)gm * N + gn] : 0.f;
    }
  }
  for (int k0 = 0; k0 < K; k0 += kBlock) {
    stage_block(lhs, lhs_scales, M, K, m0, k0, a_tile);
    stage_block(rhs, rhs_scales, N, K, n0, k0, b_tile);
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < kBlock; ++kk) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = a_tile[kk][ty + 16 * i];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = b_tile[kk][tx + 16 * j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = __fadd_rn(acc[i][j], __fmul_rn(a[i], b[j]));
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
      if (gn < N) out[(size_t)gm * N + gn] = acc[i][j];
    }
  }
}

// ---- MXFP4 -> bf16 operand expansion (first dispatch of the tensor-core MX path) ------------
//
// out[r, k] = bf16(f32(f4[r, k]) * 2^(scale[r, k / 32] - 127)). The conversion is EXACT for
// every (code, scale) pair: an f4E2M1FN value has at most two significand bits and bf16 has
// the exponent range of f32, including its subnormals (the smallest product, 0.5 * 2^-127 =
// 2^-128, is f32 bit 21), so the low 16 bits of the f32 product are always zero and the
// truncation below is the identity; overflow (>= 2^128) is +-inf and a 0xFF scale is NaN in
// both formats. The ops layer then runs the validated bf16 tcgen05 NT GEMM on the expanded
// operands (f32 accumulate): bit-identical to the reference whenever the partial sums are
// exact — which they are on the reference generator's inputs (values <= 6 * 8, granularity
// 0.25) — and within f32 summation error otherwise, the bf16 GEMM's own bar.
//
// HBM-bound: reads 17/32 byte and writes 2 bytes per element. One thread expands 8 elements
// (4 packed bytes -> one 16-byte store).
//
// ABI: bindings packed [rows, K/2] u8, scales [rows, K/32] u8 (read-only), out [rows, K] bf16
// (16-byte aligned: the ops layer owns this buffer); constants rows, K (K % 32 == 0).
// workgroup_size = [256,1,1], workgroup count = [ceil(rows * K / 8 / 256), 1, 1].
extern "C" __global__ void __launch_bounds__(256)
    iree_b200_matmul_mxfp4_dequant_bf16(const uint8_t* __restrict__ packed,
                                        const uint8_t* __restrict__ scales,
                                        uint16_t* __restrict__ out, int rows, int K) {
  const int parts = K / 8;
  const long long gid = (long long)blockIdx.x * 256 + threadIdx.x;
  if (gid >= (long long)rows * parts) return;
  const int row = (int)(gid / parts), k = (int)(gid % parts) * 8;
  const float scale = f8e8m0fnu_to_f32(scales[(size_t)row * (K / kBlock) + k / kBlock]);
  const uint8_t* p = packed + ((size_t)row * K + k) / 2;
  const uint32_t word =
      (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
  uint32_t pairs[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const uint32_t lo = __float_as_uint(__fmul_rn(f4e2m1fn_to_f32((word >> (8 * i)) & 0xFu), scale));
    const uint32_t hi =
        __float_as_uint(__fmul_rn(f4e2m1fn_to_f32((word >> (8 * i + 4)) & 0xFu), scale));
    pairs[i] = (lo >> 16) | (hi & 0xFFFF0000u);
  }
  // One 16-byte store: the ops layer guarantees a 16-byte aligned |out| and K % 32 == 0.
  struct alignas(16) Quad {
    uint32_t x, y, z, w;
  };
  *reinterpret_cast<Quad*>(out + (size_t)row * K + k) = Quad{pairs[0], pairs[1], pairs[2], pairs[3]};
}
