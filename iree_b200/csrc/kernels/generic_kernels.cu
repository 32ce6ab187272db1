// generic_kernels.cu — shape-agnostic SIMT matmul for dispatches the TMA path cannot take.
//
// TMA needs 16-byte aligned base pointers and row pitches; IREE binding pointers are only
// element-aligned (samples/custom_dispatch/cuda/kernels/kernels.cu:28-34) and the
// reference's e2e shape sets contain pitches such as 330 or 1 element
// (tests/e2e/matmul/generate_e2e_matmul_tests.py:62-83). The match-and-replace guard sends
// those dispatches here instead of to the tcgen05 kernels. This is a correctness path, not
// a throughput path: 64x64 output tile per 256-thread block staged through shared memory,
// k accumulated in ascending order in f32 / i32 exactly as the reference loops do
// (tools/testing/e2e/iree-e2e-matmul-test.cc:28-49,73-86,108-121); bf16 and f16 products
// are exact in f32, so the result is bit-identical to the reference for any input.
//
// ABI (kernels.cu:7-34): bindings lhs [batch,M,K], rhs [batch,K,N] (or [batch,N,K] when
// flags bit 1 is set), out [batch,M,N]; constants M, N, K, batch, flags
// (bit 0: accumulate into out; bit 1: transpose_rhs; bit 2: truncate the result to the
// operand type, f16/bf16 only). workgroup_size = [256,1,1],
// workgroup count = [ceil(N/64), ceil(M/64), batch].

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>

namespace {

constexpr int kTile = 64;
constexpr int kTileK = 16;

__device__ __forceinline__ float to_acc(__nv_bfloat16 v) { return __bfloat162float(v); }
__device__ __forceinline__ float to_acc(__half v) { return __half2float(v); }
__device__ __forceinline__ int32_t to_acc(int8_t v) { return (int32_t)v; }

// OCP 8-bit floats as the reference decodes them (base/internal/math.h:300-362,550-560):
// f8E5M2 is IEEE-like (it is the top byte of an f16), f8E4M3FN has no infinities and one
// NaN mantissa (S.1111.111).
struct F8E4M3FN {
  uint8_t bits;
};
struct F8E5M2 {
  uint8_t bits;
};
__device__ __forceinline__ float to_acc(F8E5M2 v) {
  return __half2float(__ushort_as_half((unsigned short)((unsigned)v.bits << 8)));
}
__device__ __forceinline__ float to_acc(F8E4M3FN v) {
  const uint32_t sign = (uint32_t)(v.bits & 0x80u) << 24;
  const uint32_t e = (v.bits >> 3) & 0xfu, m = v.bits & 0x7u;
  if (e == 0) return __uint_as_float(sign | __float_as_uint((float)m * 0.001953125f));  // m * 2^-9
  if (e == 15 && m == 7) return __uint_as_float(sign | 0x7fc00000u);
  return __uint_as_float(sign | ((e + 120u) << 23) | (m << 20));
}
template <typename T> __device__ __forceinline__ T from_acc(float v);
template <> __device__ __forceinline__ __nv_bfloat16 from_acc<__nv_bfloat16>(float v) {
  return __float2bfloat16_rn(v);
}
template <> __device__ __forceinline__ __half from_acc<__half>(float v) { return __float2half_rn(v); }

template <typename TIn, typename TAcc>
__device__ __forceinline__ void matmul_generic_body(const TIn* __restrict__ lhs,
                                                    const TIn* __restrict__ rhs,
                                                    TAcc* __restrict__ out, int M, int N, int K,
                                                    int flags,
                                                    const TAcc* __restrict__ bias = nullptr) {
  __shared__ TAcc a_tile[kTileK][kTile + 1];
  __shared__ TAcc b_tile[kTileK][kTile + 1];

  const bool accumulate = (flags & 1) != 0;
  const bool transpose_rhs = (flags & 2) != 0;
  const int b = blockIdx.z;
  const int m0 = blockIdx.y * kTile;
  const int n0 = blockIdx.x * kTile;
  const int tx = threadIdx.x % 16;
  const int ty = threadIdx.x / 16;
  lhs += (size_t)b * M * K;
  rhs += (size_t)b * K * N;
  out += (size_t)b * M * N;

  // The reference seeds the k loop with the accumulator operand and then adds the
  // products in ascending k (iree-e2e-matmul-test.cc:76-84); do the same so the
  // accumulate form is bit-identical too.
  TAcc acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int gm = m0 + ty + 16 * i, gn = n0 + tx + 16 * j;
      acc[i][j] = (accumulate && gm < M && gn < N) ? out[(size_t)gm * N + gn] : TAcc(0);
    }
  }

  for (int k0 = 0; k0 < K; k0 += kTileK) {
    // A tile: consecutive threads walk k (contiguous in memory).
    for (int idx = threadIdx.x; idx < kTile * kTileK; idx += 256) {
      const int kk = idx % kTileK, mm = idx / kTileK;
      const int gm = m0 + mm, gk = k0 + kk;
      a_tile[kk][mm] = (gm < M && gk < K) ? to_acc(lhs[(size_t)gm * K + gk]) : TAcc(0);
    }
    if (transpose_rhs) {
      for (int idx = threadIdx.x; idx < kTile * kTileK; idx += 256) {
        const int kk = idx % kTileK, nn = idx / kTileK;
        const int gn = n0 + nn, gk = k0 + kk;
        b_tile[kk][nn] = (gn < N && gk < K) ? to_acc(rhs[(size_t)gn * K + gk]) : TAcc(0);
      }
    } else {
      for (int idx = threadIdx.x; idx < kTile * kTileK; idx += 256) {
        const int nn = idx % kTile, kk = idx / kTile;
        const int gn = n0 + nn, gk = k0 + kk;
        b_tile[kk][nn] = (gn < N && gk < K) ? to_acc(rhs[(size_t)gk * N + gn]) : TAcc(0);
      }
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < kTileK; ++kk) {
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
      TAcc x = acc[i][j];
      if (bias) {  // fused consumer: x + bias[n], then (flags & 8) max(x, 0), NaN propagating
        x += bias[gn];
        if (flags & 8) x = (x > TAcc(0) || x != x) ? x : TAcc(0);
      }
      if constexpr (sizeof(TIn) == 2) {
        if (flags & 4) {  // fused arith.truncf: the result buffer holds operand-type elements
          reinterpret_cast<TIn*>(out - (size_t)b * M * N)[((size_t)b * M + gm) * N + gn] =
              from_acc<TIn>(x);
          continue;
        }
      }
      out[(size_t)gm * N + gn] = x;
    }
  }
}

}  // namespace

extern "C" __global__ void __launch_bounds__(256)
    iree_b200_matmul_generic_bf16_f32(const __nv_bfloat16* __restrict__ lhs,
                                      const __nv_bfloat16* __restrict__ rhs,
                                      float* __restrict__ out, int M, int N, int K, int batch,
                                      int flags) {
  (void)batch;
  matmul_generic_body<__nv_bfloat16, float>(lhs, rhs, out, M, N, K, flags);
}

extern "C" __global__ void __launch_bounds__(256)
    iree_b200_matmul_generic_f16_f32(const __half* __restrict__ lhs, const __half* __restrict__ rhs,
                                     float* __restrict__ out, int M, int N, int K, int batch,
                                     int flags) {
  (void)batch;
  matmul_generic_body<__half, float>(lhs, rhs, out, M, N, K, flags);
}

extern "C" __global__ void __launch_bounds__(256)
    iree_b200_matmul_generic_i8_i32(const int8_t* __restrict__ lhs, const int8_t* __restrict__ rhs,
                                    int32_t* __restrict__ out, int M, int N, int K, int batch,
                                    int flags) {
  (void)batch;
  matmul_generic_body<int8_t, int32_t>(lhs, rhs, out, M, N, K, flags);
}

// Bias / activation forms (bindings lhs, rhs, bias[N], out; flags bit 3 = relu): the
// fallback of gemm_bias_kernels.cu for shapes that fail the TMA guards.
extern "C" __global__ void __launch_bounds__(256)
    iree_b200_matmul_generic_bias_bf16_f32(const __nv_bfloat16* __restrict__ lhs,
                                           const __nv_bfloat16* __restrict__ rhs,
                                           const float* __restrict__ bias, float* __restrict__ out,
                                           int M, int N, int K, int batch, int flags) {
  (void)batch;
  matmul_generic_body<__nv_bfloat16, float>(lhs, rhs, out, M, N, K, flags, bias);
}

extern "C" __global__ void __launch_bounds__(256)
    iree_b200_matmul_generic_bias_f16_f32(const __half* __restrict__ lhs,
                                          const __half* __restrict__ rhs,
                                          const float* __restrict__ bias, float* __restrict__ out,
                                          int M, int N, int K, int batch, int flags) {
  (void)batch;
  matmul_generic_body<__half, float>(lhs, rhs, out, M, N, K, flags, bias);
}

extern "C" __global__ void __launch_bounds__(256)
    iree_b200_matmul_generic_bias_i8_i32(const int8_t* __restrict__ lhs,
                                         const int8_t* __restrict__ rhs,
                                         const int32_t* __restrict__ bias,
                                         int32_t* __restrict__ out, int M, int N, int K, int batch,
                                         int flags) {
  (void)batch;
  matmul_generic_body<int8_t, int32_t>(lhs, rhs, out, M, N, K, flags, bias);
}

extern "C" __global__ void __launch_bounds__(256)
    iree_b200_matmul_generic_f8e4m3fn_f32(const F8E4M3FN* __restrict__ lhs,
                                          const F8E4M3FN* __restrict__ rhs,
                                          float* __restrict__ out, int M, int N, int K, int batch,
                                          int flags) {
  (void)batch;
  matmul_generic_body<F8E4M3FN, float>(lhs, rhs, out, M, N, K, flags);
}

extern "C" __global__ void __launch_bounds__(256)
    iree_b200_matmul_generic_f8e5m2_f32(const F8E5M2* __restrict__ lhs,
                                        const F8E5M2* __restrict__ rhs, float* __restrict__ out,
                                        int M, int N, int K, int batch, int flags) {
  (void)batch;
  matmul_generic_body<F8E5M2, float>(lhs, rhs, out, M, N, K, flags);
}
