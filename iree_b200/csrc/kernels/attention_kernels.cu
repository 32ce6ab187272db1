// attention_kernels.cu — extern "C" entry points for iree_linalg_ext.attention.
//
// ABI (reference: samples/custom_dispatch/cuda/kernels/kernels.cu:7-34): one pointer per
// binding in layout order, then one 32-bit value per push constant; TMA-fed kernels take
// their CUtensorMaps after the ABI parameters (iree_b200_tma_recipe.h).
//
//   iree_b200_attention_prefill_{f16,bf16}   tcgen05, K1 == N == 128, no mask operand
//     bindings  Q [B,M,128] ro, K [B,K2,128] ro, V [B,K2,128] ro, O [B,M,128]
//     constants B, M, K2, scale (f32 bits), flags      grid [ceil(M/256), B, 1]  block 384
//     flags bit 0: causal mask computed in the kernel (k2 <= m + K2 - M), fully masked K2
//     blocks skipped; the grid is then [ceil(M/256) * B, 1, 1]
//   iree_b200_attention_prefill_mask_{f16,bf16}   the same with an i1 mask operand
//     bindings  Q ro, K ro, V ro, mask [B,M,K2] u8 ro, O
//     constants B, M, K2, scale, flags                 grid [ceil(M/256), B, 1]  block 384
//   iree_b200_attention_decode_split_{f16,bf16}   HBM-streaming split-KV, K1 == N == 128
//     bindings  Q ro, K ro, V ro, workspace [B,M,splits,130] f32
//     constants B, M, K2, scale, splits                grid [splits, M, B]       block 256
//   iree_b200_attention_decode_merge_{f16,bf16}
//     bindings  workspace ro, O [B,M,128]
//     constants B, M, splits                           grid [M, B, 1]            block 128
//   iree_b200_attention_decode_cluster_{f16,bf16}   single dispatch: cluster of 8 CTAs per
//     (batch, m) row, bulk-copy ring, merge over distributed shared memory
//     bindings  Q ro, K ro, V ro, O [B,M,128]
//     constants B, M, K2, scale, flags                 grid [8, M, B] (cluster 8x1x1)  block 288
//   iree_b200_attention_generic_{f16,bf16,f32}   any shape within the guards, optional mask
//     bindings  Q ro, K ro, V ro, mask ro (any valid pointer when unused), O
//     constants B, M, K1, K2, N, scale, mask_kind (0 none, 1 i1-as-u8, 2 f32 additive,
//               3 causal computed in the kernel: no operand)
//                                                      grid [M, B, 1]            block 128
// Semantics for all of them: the oracle's (oracle/oracle.c), i.e. the llvm-cpu lowering.

#include "attention_sm100.cuh"

using namespace b200;

//===----------------------------------------------------------------------===//
// Prefill (tcgen05)
//===----------------------------------------------------------------------===//

using PrefillCfg = AttnPrefillConfig<128>;

#define B200_PREFILL(NAME, BF16, POLY)                                                         \
  extern "C" __device__ const iree_b200_tma_recipe_t NAME##__tma =                             \
      make_attention_prefill_recipe<PrefillCfg, BF16>();                                       \
  extern "C" __global__ void __launch_bounds__(PrefillCfg::kThreads, 1)                        \
      NAME(const void* __restrict__ q, const void* __restrict__ k, const void* __restrict__ v, \
           void* __restrict__ out, int B, int M, int K2, int scale_bits, int flags,            \
           const __grid_constant__ TensorMap tm_q, const __grid_constant__ TensorMap tm_k,     \
           const __grid_constant__ TensorMap tm_v) {                                           \
    (void)q;                                                                                   \
    (void)k;                                                                                   \
    (void)v;                                                                                   \
    attention_prefill_body<PrefillCfg, BF16, POLY>(out, B, M, K2, scale_bits, flags, &tm_q,    \
                                                   &tm_k, &tm_v);                              \
  }

B200_PREFILL(iree_b200_attention_prefill_f16, false, 0)
B200_PREFILL(iree_b200_attention_prefill_bf16, true, 0)
// Same kernel with 1 / 2 of every 4 exp2 evaluated by poly_exp2 on the FMA pipe.
B200_PREFILL(iree_b200_attention_prefill_f16_p25, false, 1)
B200_PREFILL(iree_b200_attention_prefill_bf16_p25, true, 1)
B200_PREFILL(iree_b200_attention_prefill_f16_p50, false, 2)
B200_PREFILL(iree_b200_attention_prefill_bf16_p50, true, 2)

// Same kernel, launchable by an unmodified IREE CUDA driver: no tensor-map parameters; warp 3
// patches the host-encoded templates of the twin export (tensormap_templates.inc) with this
// dispatch's pointers and extents — Q [B, M, D], K and V [B, K2, D] — into the slots of the SM
// it runs on (common.cuh, tensormap_build_rank3). flags bit 0 = causal, as in the twin.
#define B200_TENSORMAP_TEMPLATE(NAME, ...) \
  __device__ __align__(128) const uint64_t NAME##__tmt[3][16] = {__VA_ARGS__};
#include "tensormap_templates.inc"
#undef B200_TENSORMAP_TEMPLATE
__device__ __align__(128) TensorMap iree_b200_attention_dtm_slots[kTensorMapSlotSms][4];
static_assert(PrefillCfg::kSmemBytes > 232448 / 2 && PrefillCfg::kTmemCols == 512,
              "per-SM tensor-map slots need one resident CTA per SM");

#define B200_PREFILL_DTM(NAME, TEMPLATE_OF, BF16)                                              \
  extern "C" __global__ void __launch_bounds__(PrefillCfg::kThreads, 1)                        \
      NAME(const void* __restrict__ q, const void* __restrict__ k, const void* __restrict__ v, \
           void* __restrict__ out, int B, int M, int K2, int scale_bits, int flags) {          \
    __shared__ __align__(128) TensorMap staging[3];                                            \
    TensorMap* slots = iree_b200_attention_dtm_slots[tensormap_slot_index()];                  \
    if ((threadIdx.x >> 5) == 3) {                                                             \
      tensormap_build_rank3(slots + 0, staging + 0, TEMPLATE_OF##__tmt[0], q,                  \
                            PrefillCfg::kHeadDim, M, B, 2);                                    \
      tensormap_build_rank3(slots + 1, staging + 1, TEMPLATE_OF##__tmt[1], k,                  \
                            PrefillCfg::kHeadDim, K2, B, 2);                                   \
      tensormap_build_rank3(slots + 2, staging + 2, TEMPLATE_OF##__tmt[2], v,                  \
                            PrefillCfg::kHeadDim, K2, B, 2);                                   \
    }                                                                                          \
    attention_prefill_body<PrefillCfg, BF16, 0, false, false, true>(                           \
        out, B, M, K2, scale_bits, flags, slots, slots + 1, slots + 2);                        \
  }

B200_PREFILL_DTM(iree_b200_attention_prefill_f16_dtm, iree_b200_attention_prefill_f16, false)
B200_PREFILL_DTM(iree_b200_attention_prefill_bf16_dtm, iree_b200_attention_prefill_bf16, true)

#define B200_PREFILL_MASK(NAME, BF16)                                                          \
  extern "C" __device__ const iree_b200_tma_recipe_t NAME##__tma =                             \
      make_attention_prefill_recipe<PrefillCfg, BF16>();                                       \
  extern "C" __global__ void __launch_bounds__(PrefillCfg::kThreads, 1)                        \
      NAME(const void* __restrict__ q, const void* __restrict__ k, const void* __restrict__ v, \
           const void* __restrict__ mask, void* __restrict__ out, int B, int M, int K2,        \
           int scale_bits, int flags, const __grid_constant__ TensorMap tm_q,                  \
           const __grid_constant__ TensorMap tm_k, const __grid_constant__ TensorMap tm_v) {   \
    (void)q;                                                                                   \
    (void)k;                                                                                   \
    (void)v;                                                                                   \
    attention_prefill_body<PrefillCfg, BF16, 0, false, true>(                                  \
        out, B, M, K2, scale_bits, flags, &tm_q, &tm_k, &tm_v, nullptr, 0, 1,                  \
        static_cast<const uint8_t*>(mask));                                                    \
  }

B200_PREFILL_MASK(iree_b200_attention_prefill_mask_f16, false)
B200_PREFILL_MASK(iree_b200_attention_prefill_mask_bf16, true)

// Tail-split form (attention_sm100.cuh, kSplit): 1-D grid over units, the units of the last,
// partly filled wave cut into |split| CTAs each, plus the merge of their records.
//   iree_b200_attention_prefill_split_{f16,bf16}
//     bindings  Q ro, K ro, V ro, O, workspace [tail*split][256*(128+2)] f32
//     constants B, M, K2, scale, full_units, split     grid [full_units + tail*split, 1, 1]
//   iree_b200_attention_prefill_merge_{f16,bf16}
//     bindings  workspace ro, O
//     constants B, M, full_units, split, tail_units    grid [32, tail_units, 1]  block 256
#define B200_PREFILL_SPLIT(NAME, BF16)                                                         \
  extern "C" __device__ const iree_b200_tma_recipe_t NAME##__tma =                             \
      make_attention_prefill_recipe<PrefillCfg, BF16>();                                       \
  extern "C" __global__ void __launch_bounds__(PrefillCfg::kThreads, 1)                        \
      NAME(const void* __restrict__ q, const void* __restrict__ k, const void* __restrict__ v, \
           void* __restrict__ out, float* __restrict__ workspace, int B, int M, int K2,        \
           int scale_bits, int full_units, int split, const __grid_constant__ TensorMap tm_q,  \
           const __grid_constant__ TensorMap tm_k, const __grid_constant__ TensorMap tm_v) {   \
    (void)q;                                                                                   \
    (void)k;                                                                                   \
    (void)v;                                                                                   \
    attention_prefill_body<PrefillCfg, BF16, 0, true>(out, B, M, K2, scale_bits, 0, &tm_q,     \
                                                      &tm_k, &tm_v, workspace, full_units,     \
                                                      split);                                  \
  }

B200_PREFILL_SPLIT(iree_b200_attention_prefill_split_f16, false)
B200_PREFILL_SPLIT(iree_b200_attention_prefill_split_bf16, true)

namespace {

// One CTA merges 8 rows of one split unit: thread -> (row, 4 columns).
template <bool kBf16>
__device__ __forceinline__ void attention_prefill_merge_body(const float* __restrict__ workspace,
                                                             uint16_t* __restrict__ out, int M,
                                                             int full_units, int split) {
  constexpr int D = PrefillCfg::kHeadDim;
  constexpr int kRows = PrefillCfg::kTilesPerCta * PrefillCfg::kBlockM;
  constexpr int kRecord = kRows * (D + 2);
  const int tail_unit = blockIdx.y;
  const int unit = full_units + tail_unit;
  const int tiles_m = (M + kRows - 1) / kRows;
  const int batch = unit / tiles_m;
  const int r = blockIdx.x * 8 + (int)(threadIdx.x >> 5);
  const int gm = (unit % tiles_m) * kRows + r;
  const int col = (int)(threadIdx.x & 31) * 4;
  if (gm >= M) return;
  const float* rec0 = workspace + (size_t)tail_unit * split * kRecord;
  float gmax = -FLT_MAX;
  for (int p = 0; p < split; ++p) gmax = fmaxf(gmax, rec0[(size_t)p * kRecord + kRows * D + r]);
  float sum = 0.f;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int p = 0; p < split; ++p) {
    const float* rec = rec0 + (size_t)p * kRecord;
    const float w = exp2f(rec[kRows * D + r] - gmax);
    sum = fmaf(w, rec[kRows * (D + 1) + r], sum);
    const float4 a = *reinterpret_cast<const float4*>(rec + (size_t)r * D + col);
    acc.x = fmaf(w, a.x, acc.x);
    acc.y = fmaf(w, a.y, acc.y);
    acc.z = fmaf(w, a.z, acc.z);
    acc.w = fmaf(w, a.w, acc.w);
  }
  const float rcp = 1.0f / sum;  // TileAttention.cpp:141-146
  uint2 o;
  o.x = pack2<kBf16>(acc.x * rcp, acc.y * rcp);
  o.y = pack2<kBf16>(acc.z * rcp, acc.w * rcp);
  *reinterpret_cast<uint2*>(out + ((size_t)batch * M + gm) * D + col) = o;
}

}  // namespace

#define B200_PREFILL_MERGE(SUFFIX, BF16)                                                       \
  extern "C" __global__ void __launch_bounds__(256) iree_b200_attention_prefill_merge_##SUFFIX( \
      const float* __restrict__ workspace, uint16_t* __restrict__ out, int B, int M,           \
      int full_units, int split, int tail_units) {                                             \
    (void)B;                                                                                   \
    (void)tail_units;                                                                          \
    attention_prefill_merge_body<BF16>(workspace, out, M, full_units, split);                  \
  }

B200_PREFILL_MERGE(f16, false)
B200_PREFILL_MERGE(bf16, true)

//===----------------------------------------------------------------------===//
// Decode: split-KV streaming kernel + merge. HBM bound: every K and V byte is read once.
//===----------------------------------------------------------------------===//

// The device bodies (also instantiated by no_alloc/attention_decode_no_alloc.cu).
#include "attention_decode_split.cuh"

#define B200_DECODE(SUFFIX, BF16)                                                              \
  extern "C" __global__ void __launch_bounds__(AttnDecodeConfig::kThreads)                     \
      iree_b200_attention_decode_split_##SUFFIX(                                               \
          const uint16_t* __restrict__ q, const uint16_t* __restrict__ k,                      \
          const uint16_t* __restrict__ v, float* __restrict__ workspace, int B, int M, int K2, \
          int scale_bits, int splits) {                                                        \
    attention_decode_split_body<BF16>(q, k, v, workspace, B, M, K2, scale_bits, splits);       \
  }                                                                                            \
  extern "C" __global__ void __launch_bounds__(128) iree_b200_attention_decode_merge_##SUFFIX( \
      const float* __restrict__ workspace, uint16_t* __restrict__ out, int B, int M,           \
      int splits) {                                                                            \
    attention_decode_merge_body<BF16>(workspace, out, B, M, splits);                           \
  }

B200_DECODE(f16, false)
B200_DECODE(bf16, true)

//===----------------------------------------------------------------------===//
// Decode, single dispatch. One cluster of 8 CTAs per (batch, m) row. Inside a CTA a producer
// warp keeps a 6-stage ring of [32 K rows | 32 V rows] filled with 1-D bulk copies (80 KiB
// in flight per CTA, independent of what the consumers are doing); 16 half-warps consume
// two rows per stage each, every half-warp carrying its own online-softmax state (max, sum,
// 128 sums spread over 16 lanes x 8 columns), so the streaming loop has no block-wide
// synchronisation at all. The 16 states merge in shared memory, the 8 CTA records are pushed
// into CTA 0 with st.shared::cluster and merged there. Same arithmetic as the split + merge pair.
//===----------------------------------------------------------------------===//

namespace {

template <bool kBf16>
__device__ __forceinline__ void attention_decode_cluster_body(const uint16_t* __restrict__ q,
                                                              const uint16_t* __restrict__ k,
                                                              const uint16_t* __restrict__ v,
                                                              uint16_t* __restrict__ out, int B,
                                                              int M, int K2, int scale_bits) {
  using Cfg = AttnDecodeClusterConfig;
  constexpr int D = Cfg::kHeadDim;
  constexpr int S = Cfg::kStages;
  constexpr int R = Cfg::kStageRows;
  constexpr int RF = Cfg::kRecordFloats;
  extern __shared__ uint8_t smem_raw[];
  (void)B;

  const uint32_t smem_base = (smem_u32(smem_raw) + 127u) & ~127u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  float* group_rec = reinterpret_cast<float*>(smem_gen + S * Cfg::kStageBytes);
  float* cta_rec = group_rec + Cfg::kGroups * RF;
  const uint32_t cta_rec_addr = smem_base + S * Cfg::kStageBytes + Cfg::kGroups * RF * 4;
  const uint32_t bar_full = (cta_rec_addr + Cfg::kSplits * RF * 4 + 7u) & ~7u;
  const uint32_t bar_empty = bar_full + S * 8;

  const int split = (int)cluster_ctarank(), m = blockIdx.y, b = blockIdx.z;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int rows_per_split = (K2 + Cfg::kSplits - 1) / Cfg::kSplits;
  const int row0 = min(K2, split * rows_per_split);
  const int row_end = min(K2, row0 + rows_per_split);
  const int num_stages = (row_end - row0 + R - 1) / R;

  if (tid == 0) {
    for (int s = 0; s < S; ++s) {
      mbar_init(bar_full + 8 * s, 1);
      mbar_init(bar_empty + 8 * s, Cfg::kConsumerWarps);
    }
    fence_mbar_init();
  }
  __syncthreads();

  if (warp == Cfg::kConsumerWarps) {
    // ---- producer ----
    if (lane == 0) {
      const uint8_t* kb = reinterpret_cast<const uint8_t*>(k) + ((size_t)b * K2 + row0) * Cfg::kRowBytes;
      const uint8_t* vb = reinterpret_cast<const uint8_t*>(v) + ((size_t)b * K2 + row0) * Cfg::kRowBytes;
      for (int it = 0; it < num_stages; ++it) {
        const int s = it % S;
        if (it >= S) mbar_wait(bar_empty + 8 * s, ((it / S) - 1) & 1);
        const int rows = min(R, row_end - row0 - it * R);
        const uint32_t bytes = (uint32_t)rows * Cfg::kRowBytes;
        const uint32_t dst = smem_base + s * Cfg::kStageBytes;
        mbar_expect_tx(bar_full + 8 * s, 2 * bytes);
        bulk_load_1d(dst, kb + (size_t)it * R * Cfg::kRowBytes, bytes, bar_full + 8 * s);
        bulk_load_1d(dst + R * Cfg::kRowBytes, vb + (size_t)it * R * Cfg::kRowBytes, bytes,
                     bar_full + 8 * s);
      }
    }
  } else {
    // ---- consumers ----
    const int sub = lane & 15;             // 16-byte column slab of the 256-byte row
    const int grp = warp * 2 + (lane >> 4);  // half-warp id, owns rows 2*grp, 2*grp+1 of a stage
    float qf[8];
    {
      // Q' = round_T(s' * q) (AggregatedOpInterfaceImpl.cpp:355-358).
      const float mult = attention_q_multiplier<kBf16>(__int_as_float(scale_bits));
      const uint4 qv = *reinterpret_cast<const uint4*>(q + ((size_t)b * M + m) * D + sub * 8);
      unpack8<kBf16>(qv, qf);
#pragma unroll
      for (int e = 0; e < 8; ++e) qf[e] = elem_to_f32<kBf16>(f32_to_elem<kBf16>(qf[e] * mult));
    }
    float run_max = -FLT_MAX, run_sum = 0.f;  // TileAttention.cpp:78-96
    float acc[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = 0.f;

    for (int it = 0; it < num_stages; ++it) {
      const int s = it % S;
      mbar_wait(bar_full + 8 * s, (it / S) & 1);
      const uint8_t* stage = smem_gen + s * Cfg::kStageBytes;
      const int rows = min(R, row_end - row0 - it * R);
      const int r0 = grp * 2;
      const bool ok0 = r0 < rows, ok1 = r0 + 1 < rows;
      const uint4 k0 = *reinterpret_cast<const uint4*>(stage + r0 * Cfg::kRowBytes + sub * 16);
      const uint4 k1 = *reinterpret_cast<const uint4*>(stage + (r0 + 1) * Cfg::kRowBytes + sub * 16);
      const uint4 v0 =
          *reinterpret_cast<const uint4*>(stage + (R + r0) * Cfg::kRowBytes + sub * 16);
      const uint4 v1 =
          *reinterpret_cast<const uint4*>(stage + (R + r0 + 1) * Cfg::kRowBytes + sub * 16);
      float f0[8], f1[8];
      unpack8<kBf16>(k0, f0);
      unpack8<kBf16>(k1, f1);
      float d0 = 0.f, d1 = 0.f;
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        d0 = fmaf(qf[e], f0[e], d0);
        d1 = fmaf(qf[e], f1[e], d1);
      }
#pragma unroll
      for (int o = 8; o > 0; o >>= 1) {
        d0 += __shfl_xor_sync(0xffffffffu, d0, o);
        d1 += __shfl_xor_sync(0xffffffffu, d1, o);
      }
      // All shared-memory reads of this stage are done: hand it back to the producer.
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_empty + 8 * s);
      if (ok0) {
        const float s0 = d0, s1 = ok1 ? d1 : -INFINITY;
        const float new_max = fmaxf(run_max, fmaxf(s0, s1));
        const float norm = fast_exp2(run_max - new_max);
        const float p0 = fast_exp2(s0 - new_max), p1 = fast_exp2(s1 - new_max);
        run_sum = fmaf(norm, run_sum, p0 + p1);
        run_max = new_max;
        // P is truncated to V's element type before P.V (:613-618).
        const float t0 = elem_to_f32<kBf16>(f32_to_elem<kBf16>(p0));
        const float t1 = elem_to_f32<kBf16>(f32_to_elem<kBf16>(p1));
        unpack8<kBf16>(v0, f0);
        unpack8<kBf16>(v1, f1);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          float a = fmaf(t0, f0[e], norm * acc[e]);
          if (ok1) a = fmaf(t1, f1[e], a);   // a stale row may hold Inf/NaN bits: skip, not x0
          acc[e] = a;
        }
      }
    }
    // ---- the 16 half-warp states -> shared memory ----
    float* rec = group_rec + grp * RF;
#pragma unroll
    for (int e = 0; e < 8; ++e) rec[sub * 8 + e] = acc[e];
    if (sub == 0) {
      rec[D] = run_max;
      rec[D + 1] = run_sum;
    }
  }
  __syncthreads();
  if (tid < D) {
    float gmax = -FLT_MAX;
#pragma unroll
    for (int g = 0; g < Cfg::kGroups; ++g) gmax = fmaxf(gmax, group_rec[g * RF + D]);
    float sum = 0.f, a = 0.f;
#pragma unroll
    for (int g = 0; g < Cfg::kGroups; ++g) {
      const float w = exp2f(group_rec[g * RF + D] - gmax);
      sum = fmaf(w, group_rec[g * RF + D + 1], sum);
      a = fmaf(w, group_rec[g * RF + tid], a);
    }
    // Push this CTA's record into CTA 0's shared memory (slot |split|): one cluster barrier
    // then orders all eight records before CTA 0 reads them locally, and the other CTAs
    // may retire right after arriving (nobody reads their shared memory).
    const uint32_t slot = map_to_cta(cta_rec_addr + (uint32_t)split * RF * 4, 0);
    st_shared_cluster_f32(slot + tid * 4, a);
    if (tid == 0) {
      st_shared_cluster_f32(slot + D * 4, gmax);
      st_shared_cluster_f32(slot + (D + 1) * 4, sum);
    }
  }
  cluster_arrive();  // .release: the stores above are visible to whoever acquires
  if (split != 0) return;
  cluster_wait();
  if (tid < D) {
    float gmax = -FLT_MAX;
#pragma unroll
    for (int c = 0; c < Cfg::kSplits; ++c) gmax = fmaxf(gmax, cta_rec[c * RF + D]);
    float sum = 0.f, a = 0.f;
#pragma unroll
    for (int c = 0; c < Cfg::kSplits; ++c) {
      const float w = exp2f(cta_rec[c * RF + D] - gmax);
      sum = fmaf(w, cta_rec[c * RF + D + 1], sum);
      a = fmaf(w, cta_rec[c * RF + tid], a);
    }
    const float rcp = 1.0f / sum;  // TileAttention.cpp:141-146
    out[((size_t)b * M + m) * D + tid] = f32_to_elem<kBf16>(rcp * a);
  }
}

}  // namespace

#define B200_DECODE_CLUSTER(SUFFIX, BF16)                                                      \
  extern "C" __global__ void __cluster_dims__(AttnDecodeClusterConfig::kSplits, 1, 1)          \
      __launch_bounds__(AttnDecodeClusterConfig::kThreads)                                     \
          iree_b200_attention_decode_cluster_##SUFFIX(                                         \
              const uint16_t* __restrict__ q, const uint16_t* __restrict__ k,                  \
              const uint16_t* __restrict__ v, uint16_t* __restrict__ out, int B, int M,        \
              int K2, int scale_bits, int flags) {                                             \
    (void)flags;                                                                               \
    attention_decode_cluster_body<BF16>(q, k, v, out, B, M, K2, scale_bits);                   \
  }

B200_DECODE_CLUSTER(f16, false)
B200_DECODE_CLUSTER(bf16, true)

//===----------------------------------------------------------------------===//
// Generic SIMT kernel: any B/M/K1/K2/N within the guards, optional mask operand, f16 /
// bf16 / f32 elements. Correctness path (masks, odd head dims, unaligned pointers, the
// reference's f32 golden vectors): one CTA per (b, m) row, online softmax over K2 tiles of
// 128 exactly as OnlineAttentionOp::decomposeOperation.
//===----------------------------------------------------------------------===//

namespace {

enum : int { kElemF16 = 0, kElemBF16 = 1, kElemF32 = 2 };

template <int kType>
__device__ __forceinline__ float ld_elem(const void* p, size_t i) {
  if constexpr (kType == kElemF32) {
    return reinterpret_cast<const float*>(p)[i];
  } else {
    return elem_to_f32<kType == kElemBF16>(reinterpret_cast<const uint16_t*>(p)[i]);
  }
}

template <int kType>
__device__ __forceinline__ float round_elem(float v) {
  if constexpr (kType == kElemF32) {
    return v;
  } else {
    return elem_to_f32<kType == kElemBF16>(f32_to_elem<kType == kElemBF16>(v));
  }
}

template <int kType>
__device__ __forceinline__ void st_elem(void* p, size_t i, float v) {
  if constexpr (kType == kElemF32) {
    reinterpret_cast<float*>(p)[i] = v;
  } else {
    reinterpret_cast<uint16_t*>(p)[i] = f32_to_elem<kType == kElemBF16>(v);
  }
}

template <int kType>
__device__ __forceinline__ void attention_generic_body(const void* __restrict__ q,
                                                       const void* __restrict__ k,
                                                       const void* __restrict__ v,
                                                       const void* __restrict__ mask,
                                                       void* __restrict__ out, int B, int M,
                                                       int K1, int K2, int N, int scale_bits,
                                                       int mask_kind) {
  using Cfg = AttnGenericConfig;
  constexpr int T = Cfg::kThreads;
  __shared__ float q_s[Cfg::kMaxK1];
  __shared__ float p_s[T];
  __shared__ float red[T / 32];
  const int m = blockIdx.x, b = blockIdx.y, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  (void)B;
  // s' = scale*log2e in the scale's type, cast to Q's type (:336-358).
  const float mult = round_elem<kType>(round_elem<kType>(__int_as_float(scale_bits)) *
                                       round_elem<kType>(1.4426950408889634074f));
  for (int d = tid; d < K1; d += T) {
    q_s[d] = round_elem<kType>(mult * ld_elem<kType>(q, ((size_t)b * M + m) * K1 + d));
  }
  __syncthreads();
  float run_max = -FLT_MAX, run_sum = 0.f;
  float acc[Cfg::kMaxN / T];
#pragma unroll
  for (int i = 0; i < Cfg::kMaxN / T; ++i) acc[i] = 0.f;
  for (int t0 = 0; t0 < K2; t0 += T) {
    const int j = t0 + tid;
    float s = -INFINITY;
    if (j < K2) {
      const size_t kbase = ((size_t)b * K2 + j) * K1;
      float a = 0.f;
      for (int d = 0; d < K1; ++d) a = fmaf(q_s[d], ld_elem<kType>(k, kbase + d), a);
      const size_t midx = ((size_t)b * M + m) * K2 + j;
      if (mask_kind == 1) {
        a = a + (reinterpret_cast<const uint8_t*>(mask)[midx] ? 0.f : -INFINITY);
      } else if (mask_kind == 2) {
        a = a + reinterpret_cast<const float*>(mask)[midx] * 1.4426950408889634074f;
      } else if (mask_kind == 3) {  // causal, computed: k2 <= m + (K2 - M)
        a = a + (j <= m + (K2 - M) ? 0.f : -INFINITY);
      }
      s = a;
    }
    float wmx = s;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) wmx = fmaxf(wmx, __shfl_xor_sync(0xffffffffu, wmx, o));
    if (lane == 0) red[warp] = wmx;
    __syncthreads();
    float new_max = run_max;
#pragma unroll
    for (int w = 0; w < T / 32; ++w) new_max = fmaxf(new_max, red[w]);
    const float norm = exp2f(run_max - new_max);
    const float p = j < K2 ? exp2f(s - new_max) : 0.f;
    float psum = p;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) psum += __shfl_xor_sync(0xffffffffu, psum, o);
    __syncthreads();
    p_s[tid] = round_elem<kType>(p);
    if (lane == 0) red[warp] = psum;
    __syncthreads();
    float tot = 0.f;
#pragma unroll
    for (int w = 0; w < T / 32; ++w) tot += red[w];
    run_sum = norm * run_sum + tot;
    run_max = new_max;
    const int tl = min(T, K2 - t0);
#pragma unroll
    for (int i = 0; i < Cfg::kMaxN / T; ++i) {
      const int n = tid + i * T;
      if (n < N) {
        float a = norm * acc[i];
        for (int jj = 0; jj < tl; ++jj) {
          a = fmaf(p_s[jj], ld_elem<kType>(v, ((size_t)b * K2 + t0 + jj) * N + n), a);
        }
        acc[i] = a;
      }
    }
    __syncthreads();
  }
  const float rcp = 1.0f / run_sum;
#pragma unroll
  for (int i = 0; i < Cfg::kMaxN / T; ++i) {
    const int n = tid + i * T;
    if (n < N) st_elem<kType>(out, ((size_t)b * M + m) * N + n, rcp * acc[i]);
  }
}

}  // namespace

#define B200_GENERIC(SUFFIX, TYPE)                                                            \
  extern "C" __global__ void __launch_bounds__(AttnGenericConfig::kThreads)                   \
      iree_b200_attention_generic_##SUFFIX(                                                   \
          const void* __restrict__ q, const void* __restrict__ k, const void* __restrict__ v, \
          const void* __restrict__ mask, void* __restrict__ out, int B, int M, int K1, int K2, \
          int N, int scale_bits, int mask_kind) {                                             \
    attention_generic_body<TYPE>(q, k, v, mask, out, B, M, K1, K2, N, scale_bits, mask_kind); \
  }

B200_GENERIC(f16, kElemF16)
B200_GENERIC(bf16, kElemBF16)
B200_GENERIC(f32, kElemF32)
