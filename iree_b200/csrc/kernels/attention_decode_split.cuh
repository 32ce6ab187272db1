// attention_decode_split.cuh — device bodies of the split-KV decode attention pair
// (iree_b200_attention_decode_split_* / _merge_*, attention_kernels.cu): one CTA per
// (split, m, batch) streams its share of K and V once and writes a partial record
// [kHeadDim unnormalised f32 sums, running max, running sum] to the workspace; the merge
// kernel combines the records of a row. Shared by attention_kernels.cu (workspace = a
// binding) and no_alloc/attention_decode_no_alloc.cu (workspace = a static device array).
// Semantics: the oracle's (oracle/oracle.c), i.e. the llvm-cpu lowering.

#ifndef IREE_B200_KERNELS_ATTENTION_DECODE_SPLIT_CUH_
#define IREE_B200_KERNELS_ATTENTION_DECODE_SPLIT_CUH_

#include <float.h>

#include "attention_config.h"
#include "attention_sm100.cuh"

namespace b200 {

__device__ __forceinline__ uint4 ld_stream_16(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

template <bool kBf16>
__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    f[2 * e] = elem_to_f32<kBf16>((uint16_t)(w[e] & 0xffffu));
    f[2 * e + 1] = elem_to_f32<kBf16>((uint16_t)(w[e] >> 16));
  }
}

template <bool kBf16>
__device__ __forceinline__ void attention_decode_split_body(const uint16_t* __restrict__ q,
                                                            const uint16_t* __restrict__ k,
                                                            const uint16_t* __restrict__ v,
                                                            float* __restrict__ workspace, int B,
                                                            int M, int K2, int scale_bits,
                                                            int splits) {
  using Cfg = AttnDecodeConfig;
  constexpr int D = Cfg::kHeadDim;
  constexpr int R = Cfg::kChunkRows;
  __shared__ float s_s[R];            // scores, then P (already truncated to the V type)
  __shared__ float red[8];
  __shared__ float acc_s[16][D + 4];  // per row-group partial sums

  const int split = blockIdx.x, m = blockIdx.y, b = blockIdx.z;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int sub = tid & 15;   // 16-byte column slab of the 256-byte row
  const int grp = tid >> 4;   // 16 row groups
  const int rows_per_split = (K2 + splits - 1) / splits;
  const int row0 = split * rows_per_split;
  const int row_end = min(K2, row0 + rows_per_split);
  (void)B;

  // Q' slab of this thread: round_T(s' * q) (AggregatedOpInterfaceImpl.cpp:355-358).
  float qf[8];
  {
    const float mult = attention_q_multiplier<kBf16>(__int_as_float(scale_bits));
    const uint4 qv = *reinterpret_cast<const uint4*>(q + ((size_t)b * M + m) * D + sub * 8);
    unpack8<kBf16>(qv, qf);
#pragma unroll
    for (int e = 0; e < 8; ++e) qf[e] = elem_to_f32<kBf16>(f32_to_elem<kBf16>(qf[e] * mult));
  }

  float run_max = -FLT_MAX, run_sum = 0.f;
  float acc[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) acc[e] = 0.f;

  for (int c0 = row0; c0 < row_end; c0 += R) {
    const int rows = min(R, row_end - c0);
    const uint16_t* kbase = k + ((size_t)b * K2 + c0) * D + sub * 8;
    const uint16_t* vbase = v + ((size_t)b * K2 + c0) * D + sub * 8;
    // ---- scores: 16 rows per pass, 16 lanes per row ----
#pragma unroll 4
    for (int r = grp; r < R; r += 16) {
      float dot = 0.f;
      if (r < rows) {
        float kf[8];
        unpack8<kBf16>(ld_stream_16(kbase + (size_t)r * D), kf);
#pragma unroll
        for (int e = 0; e < 8; ++e) dot = fmaf(qf[e], kf[e], dot);
      }
      dot += __shfl_xor_sync(0xffffffffu, dot, 8);
      dot += __shfl_xor_sync(0xffffffffu, dot, 4);
      dot += __shfl_xor_sync(0xffffffffu, dot, 2);
      dot += __shfl_xor_sync(0xffffffffu, dot, 1);
      if (sub == 0) s_s[r] = r < rows ? dot : -INFINITY;
    }
    __syncthreads();
    // ---- block max ----
    float mx = s_s[tid];  // R == blockDim.x
    float wmx = mx;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) wmx = fmaxf(wmx, __shfl_xor_sync(0xffffffffu, wmx, o));
    if (lane == 0) red[warp] = wmx;
    __syncthreads();
    float new_max = run_max;
#pragma unroll
    for (int w = 0; w < 8; ++w) new_max = fmaxf(new_max, red[w]);
    const float norm = exp2f(run_max - new_max);
    const float p = exp2f(mx - new_max);
    float psum = p;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) psum += __shfl_xor_sync(0xffffffffu, psum, o);
    __syncthreads();  // everyone has read red[] and s_s[tid]
    s_s[tid] = elem_to_f32<kBf16>(f32_to_elem<kBf16>(p));
    if (lane == 0) red[warp] = psum;
    __syncthreads();
    float tot = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) tot += red[w];
    run_sum = norm * run_sum + tot;
    run_max = new_max;
    // ---- acc = norm*acc + P.V over this thread's row group / column slab ----
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] *= norm;
#pragma unroll 4
    for (int r = grp; r < R; r += 16) {
      if (r < rows) {
        float vf[8];
        unpack8<kBf16>(ld_stream_16(vbase + (size_t)r * D), vf);
        const float pr = s_s[r];
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] = fmaf(pr, vf[e], acc[e]);
      }
    }
    __syncthreads();  // s_s / red reused by the next chunk
  }

  // ---- reduce the 16 row groups and write the partial record ----
#pragma unroll
  for (int e = 0; e < 8; ++e) acc_s[grp][sub * 8 + e] = acc[e];
  __syncthreads();
  float* rec = workspace + (((size_t)b * M + m) * splits + split) * Cfg::kRecordFloats;
  if (tid < D) {
    float a = 0.f;
#pragma unroll
    for (int g = 0; g < 16; ++g) a += acc_s[g][tid];
    rec[tid] = a;
  }
  if (tid == 0) {
    rec[D] = run_max;
    rec[D + 1] = run_sum;
  }
}

template <bool kBf16>
__device__ __forceinline__ void attention_decode_merge_body(const float* __restrict__ workspace,
                                                            uint16_t* __restrict__ out, int B,
                                                            int M, int splits) {
  using Cfg = AttnDecodeConfig;
  constexpr int D = Cfg::kHeadDim;
  const int m = blockIdx.x, b = blockIdx.y, n = threadIdx.x;
  (void)B;
  const float* rec = workspace + ((size_t)b * M + m) * splits * Cfg::kRecordFloats;
  float gmax = -FLT_MAX;
  for (int s = 0; s < splits; ++s) gmax = fmaxf(gmax, rec[s * Cfg::kRecordFloats + D]);
  float sum = 0.f, acc = 0.f;
  for (int s = 0; s < splits; ++s) {
    const float w = exp2f(rec[s * Cfg::kRecordFloats + D] - gmax);
    sum = fmaf(w, rec[s * Cfg::kRecordFloats + D + 1], sum);
    acc = fmaf(w, rec[s * Cfg::kRecordFloats + n], acc);
  }
  const float rcp = 1.0f / sum;  // TileAttention.cpp:141-146
  out[((size_t)b * M + m) * D + n] = f32_to_elem<kBf16>(rcp * acc);
}

}  // namespace b200

#endif  // IREE_B200_KERNELS_ATTENTION_DECODE_SPLIT_CUH_
