// attention_sm100.cuh — tcgen05 prefill kernel for iree_linalg_ext.attention on sm_100a.
//
// Arithmetic = the llvm-cpu lowering of the op (the oracle, oracle/oracle.c):
//   compiler/src/iree/compiler/Dialect/LinalgExt/IR/AggregatedOpInterfaceImpl.cpp:323-412
//     Q' = round_T(round_T(scale*log2e) * Q)   (in Q's element type), S = Q'.K^T in f32
//   :533-632 online softmax in the exp2 domain, P truncated to V's type before P.V (f32 acc)
//   compiler/src/iree/compiler/Dialect/LinalgExt/Transforms/TileAttention.cpp:78-96,136-152
//     max starts at -FLT_MAX, O = cast((1/sum) * acc).
// The K2 block (128) only moves where the online rescale happens, and the rescale of the
// accumulator is applied lazily (only when a row's max grew by more than 2^8), which is
// algebraically the same result; parity is therefore by tolerance (1e-3 f16 / 1e-2 bf16).
//
// Masks (the op's optional mask operand, LinalgExtOps.td:967-1071; applied to S before the
// row max as `S + select(mask, 0, -inf)`, AggregatedOpInterfaceImpl.cpp:239-291):
//   * flags bit 0 (kAttnFlagCausal): the mask is the causal one, mask[m, k2] = k2 <= m +
//     (K2 - M), and is never materialised: the softmax thread of row m turns the columns
//     beyond its limit into -inf, and K2 blocks that are masked for every row of the CTA are
//     not loaded or multiplied at all (half the work at M == K2). The grid is then 1-D over
//     units with the heaviest units (largest m) first, so the light ones fill the tail.
//   * kMask: an i1 mask tensor [B, M, K2] (one byte per element, as IREE stores i1) is a
//     fourth read-only binding; each softmax thread reads the 128 bytes of its row and block
//     (issued before it waits for S, so the latency hides behind the MMA) and turns masked
//     columns into -inf.
//
// Layouts (row-major): Q [B,M,D], K [B,K2,D], V [B,K2,D], O [B,M,D], D = 128.
//
// One CTA (384 threads) owns two 128-row Q tiles of one batch slice:
//   warp 0       TMA producer: Q tiles once, then K_j / V_j blocks into two 2-deep rings.
//   warp 1       MMA issuer (one elected thread): S_t = Q'_t.K_j^T (SS form) into TMEM,
//                O_t += P_t.V_j (TS form: P read from TMEM, V from smem, MN-major).
//                Issue order PV0_j, S0_{j+1}, PV1_j, S1_{j+1}: the softmax of one tile
//                overlaps the MMAs of the other.
//   warp 2       TMEM allocator (512 columns: S0 | S1 | O0 | O1; P_t aliases S_t).
//   warps 4-7    softmax warpgroup of tile 0, warps 8-11 of tile 1: one thread per row
//                (tcgen05.ld 32x32b), exp2 on MUFU, P packed to 16 bit and written back with
//                tcgen05.st, lazy O correction, final normalisation and store.
// Roofline: tensor-pipe bound (4*B*M*K2*D FLOP), MUFU co-limited (B*M*K2 exp2).

#pragma once

#include <float.h>

#include "attention_config.h"
#include "common.cuh"
#include "iree_b200_tma_recipe.h"

namespace b200 {

template <bool kBf16>
B200_DEVICE float elem_to_f32(uint16_t bits) {
  if constexpr (kBf16) {
    return __uint_as_float((uint32_t)bits << 16);
  } else {
    return __half2float(__ushort_as_half(bits));
  }
}

template <bool kBf16>
B200_DEVICE uint16_t f32_to_elem(float v) {
  if constexpr (kBf16) {
    return __bfloat16_as_ushort(__float2bfloat16_rn(v));
  } else {
    return __half_as_ushort(__float2half_rn(v));
  }
}

// Packs two f32 into one 32-bit register of two 16-bit elements (lo = first).
template <bool kBf16>
B200_DEVICE uint32_t pack2(float lo, float hi) {
  uint32_t r;
  if constexpr (kBf16) {
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  } else {
    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  }
  return r;
}

// 16 read-only bytes that are touched once (the mask): no L1 allocation.
B200_DEVICE uint4 ld_nc_na_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

B200_DEVICE float max3(float a, float b, float c) {
  float r;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
  return r;
}

B200_DEVICE float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// 2^x on the FMA / ALU pipes instead of the MUFU (16 results/clk/SM, the co-limiter of the
// prefill kernel): n = round(x) by the 1.5*2^23 trick, a degree-3 minimax polynomial for
// 2^f on f = x - n in [-0.5, 0.5] (max relative error 7.5e-5, well inside the 16-bit
// rounding P gets anyway), and the exponent put back with one shift-add on the bits.
B200_DEVICE float poly_exp2(float x) {
  x = fmaxf(x, -126.0f);
  const float t = x + 12582912.0f;       // low mantissa bits of t now hold n
  const float f = x - (t - 12582912.0f);
  float p = fmaf(0.0551716648f, f, 0.2426111251f);
  p = fmaf(p, f, 0.6932609677f);
  p = fmaf(p, f, 0.9999280572f);
  return __int_as_float(__float_as_int(p) + (__float_as_int(t) << 23));
}

// s' of the reference: scale*log2e evaluated in the scale's (== element) type, then cast
// to Q's type (AggregatedOpInterfaceImpl.cpp:336-358).
template <bool kBf16>
B200_DEVICE float attention_q_multiplier(float scale) {
  const float log2e_t = elem_to_f32<kBf16>(f32_to_elem<kBf16>(1.4426950408889634074f));
  const float scale_t = elem_to_f32<kBf16>(f32_to_elem<kBf16>(scale));
  return elem_to_f32<kBf16>(f32_to_elem<kBf16>(scale_t * log2e_t));
}


// kPoly: how many of every four exp2 go to poly_exp2 instead of the MUFU (0, 1 or 2).
// kSplit: the tail-split form. The grid is 1-D over work units (a unit = two 128-row Q tiles
// of one batch slice); units < |full_units| run whole, every later unit is cut into |split|
// CTAs that each walk 1/split of the K2 blocks and leave an unnormalised record
// (acc [256,128] f32, max [256], sum [256]) in |workspace| for the merge kernel. With equal
// units the last wave of a grid is otherwise only partly full (512 units on 148 SMs: 3.46
// waves); splitting just those units makes the last wave short instead.
constexpr int kAttnFlagCausal = 1;

template <class Cfg, bool kBf16, int kPoly = 0, bool kSplit = false, bool kMask = false,
          bool kDeviceMaps = false>
__device__ __forceinline__ void attention_prefill_body(void* __restrict__ out, int B, int M, int K2,
                                                       int scale_bits, int flags,
                                                       const TensorMap* tm_q, const TensorMap* tm_k,
                                                       const TensorMap* tm_v,
                                                       float* __restrict__ workspace = nullptr,
                                                       int full_units = 0, int split = 1,
                                                       const uint8_t* __restrict__ mask = nullptr) {
  constexpr int D = Cfg::kHeadDim;
  constexpr int BM = Cfg::kBlockM;
  constexpr int BN = Cfg::kBlockN;
  constexpr uint32_t kFmt = kBf16 ? FMT_BF16 : FMT_F16;
  constexpr uint32_t kIdescS = make_idesc(CFMT_F32, kFmt, kFmt, MAJOR_K, MAJOR_K, BM, BN);
  constexpr uint32_t kIdescPV = make_idesc(CFMT_F32, kFmt, kFmt, MAJOR_K, MAJOR_MN, BM, D);
  const bool causal = !kSplit && (flags & kAttnFlagCausal) != 0;
  const int causal_shift = K2 - M;  // >= 0 (selection guard): row m sees columns <= m + shift

  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t smem_q = smem_base;                                      // [2] tiles
  const uint32_t smem_k = smem_q + Cfg::kTilesPerCta * Cfg::kTileBytes;   // [kKStages]
  const uint32_t smem_v = smem_k + Cfg::kKStages * Cfg::kTileBytes;       // [kVStages]
  const uint32_t bar_base = smem_base + Cfg::kSmemTiles;
  const uint32_t bar_q_full = bar_base;                        // [2] TMA -> softmax WG
  const uint32_t bar_q_scaled = bar_q_full + 16;               // [2] softmax WG -> MMA
  const uint32_t bar_k_full = bar_q_scaled + 16;               // [kKStages]
  const uint32_t bar_k_empty = bar_k_full + 8 * Cfg::kKStages;
  const uint32_t bar_v_full = bar_k_empty + 8 * Cfg::kKStages;
  const uint32_t bar_v_empty = bar_v_full + 8 * Cfg::kVStages;
  const uint32_t bar_s_full = bar_v_empty + 8 * Cfg::kVStages;  // [2] MMA -> softmax
  const uint32_t bar_p_full = bar_s_full + 16;                  // [2 tiles][2 halves] softmax -> MMA
  const uint32_t bar_o_done = bar_p_full + 32;                  // [1] MMA -> epilogue
  const uint32_t tmem_slot = bar_o_done + 8;

  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t lane = threadIdx.x & 31;
  int batch = blockIdx.y;
  int m_cta = blockIdx.x * (Cfg::kTilesPerCta * BM);
  uint32_t num_kv = (uint32_t)((K2 + BN - 1) / BN);  // K2 blocks this CTA walks
  uint32_t kv_begin = 0;                              // first of them
  int record = -1;                                    // workspace record (split CTAs only)
  if constexpr (kSplit) {
    const int tiles_m = (M + Cfg::kTilesPerCta * BM - 1) / (Cfg::kTilesPerCta * BM);
    int unit = blockIdx.x;
    if (unit >= full_units) {
      const int r = unit - full_units;
      unit = full_units + r / split;
      const uint32_t part = (uint32_t)(r % split);
      const uint32_t per = (num_kv + split - 1) / split;  // selection guarantees >= 1 each
      kv_begin = min(num_kv, part * per);
      num_kv = min(num_kv, kv_begin + per) - kv_begin;
      record = r;
    }
    batch = unit / tiles_m;
    m_cta = (unit % tiles_m) * (Cfg::kTilesPerCta * BM);
  } else if (causal) {
    // 1-D grid, heaviest units first: unit u = (m tile counted from the bottom, batch).
    const int tiles_m = (M + Cfg::kTilesPerCta * BM - 1) / (Cfg::kTilesPerCta * BM);
    batch = (int)blockIdx.x % B;
    m_cta = (tiles_m - 1 - (int)blockIdx.x / B) * (Cfg::kTilesPerCta * BM);
    // Blocks whose first column lies beyond the last row's limit are masked for every row.
    const int last_row = min(M, m_cta + Cfg::kTilesPerCta * BM) - 1;
    num_kv = min(num_kv, (uint32_t)((last_row + causal_shift + BN) / BN));
  }

  // kDeviceMaps: |tm_*| are global slots warp 3 filled before the call (attention_kernels.cu,
  // `*_dtm`); the producer acquires them after the setup barrier.
  if (!kDeviceMaps && warp == 0 && lane == 0) {
    tma_prefetch_desc(tm_q);
    tma_prefetch_desc(tm_k);
    tma_prefetch_desc(tm_v);
  }
  if (warp == 1 && lane == 0) {
    for (int t = 0; t < 2; ++t) {
      mbar_init(bar_q_full + 8 * t, 1);
      mbar_init(bar_q_scaled + 8 * t, 4);  // one arrive per softmax warp
      mbar_init(bar_s_full + 8 * t, 1);
      mbar_init(bar_p_full + 16 * t, 4);
      mbar_init(bar_p_full + 16 * t + 8, 4);
    }
    for (int s = 0; s < Cfg::kKStages; ++s) {
      mbar_init(bar_k_full + 8 * s, 1);
      mbar_init(bar_k_empty + 8 * s, 1);
    }
    for (int s = 0; s < Cfg::kVStages; ++s) {
      mbar_init(bar_v_full + 8 * s, 1);
      mbar_init(bar_v_empty + 8 * s, 1);
    }
    mbar_init(bar_o_done, 1);
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc<1>(tmem_slot, Cfg::kTmemCols);
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_gen + (tmem_slot - smem_base));

  // Register reallocation: the softmax warpgroups hold a whole 128-column S row per thread.
  // (setmaxnreg sits at the top of each role branch so ptxas sees one limit per region.)
  if (warp < 4) asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");
  if (warp == 0) {
    //===------------------------------------------------------------------===//
    // TMA producer
    //===------------------------------------------------------------------===//
    if (elect_one()) {
      if constexpr (kDeviceMaps) {
        tensormap_fence_acquire(tm_q);
        tensormap_fence_acquire(tm_k);
        tensormap_fence_acquire(tm_v);
      }
      for (int t = 0; t < Cfg::kTilesPerCta; ++t) {
        mbar_expect_tx(bar_q_full + 8 * t, Cfg::kTileBytes);
        for (int c = 0; c < Cfg::kChunks; ++c) {
          tma_load_3d<1>(smem_q + t * Cfg::kTileBytes + c * Cfg::kChunkBytes, tm_q,
                         bar_q_full + 8 * t, c * Cfg::kChunkElems, m_cta + t * BM, batch);
        }
      }
      uint32_t ks = 0, kphase = 0, vs = 0, vphase = 0;
      for (uint32_t j = 0; j < num_kv; ++j) {
        const int row0 = (int)((kv_begin + j) * BN);
        mbar_wait(bar_k_empty + 8 * ks, kphase ^ 1);
        mbar_expect_tx(bar_k_full + 8 * ks, Cfg::kTileBytes);
        for (int c = 0; c < Cfg::kChunks; ++c) {
          tma_load_3d<1>(smem_k + ks * Cfg::kTileBytes + c * Cfg::kChunkBytes, tm_k,
                         bar_k_full + 8 * ks, c * Cfg::kChunkElems, row0, batch);
        }
        if (++ks == Cfg::kKStages) {
          ks = 0;
          kphase ^= 1;
        }
        mbar_wait(bar_v_empty + 8 * vs, vphase ^ 1);
        mbar_expect_tx(bar_v_full + 8 * vs, Cfg::kTileBytes);
        for (int c = 0; c < Cfg::kChunks; ++c) {
          // MN-major operand: one box per 64-column slab of the head dimension.
          tma_load_3d<1>(smem_v + vs * Cfg::kTileBytes + c * Cfg::kChunkBytes, tm_v,
                         bar_v_full + 8 * vs, c * Cfg::kChunkElems, row0, batch);
        }
        if (++vs == Cfg::kVStages) {
          vs = 0;
          vphase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    //===------------------------------------------------------------------===//
    // MMA issuer
    //===------------------------------------------------------------------===//
    if (elect_one()) {
      // S_t (128 f32 columns) at column t*128; O_t at 256 + t*128; P_t aliases S_t.
      // Descriptors differ only in the start-address field (low word, 16-byte units), so
      // each MMA's descriptor is base + immediate: nothing to keep live across the loop.
      const uint64_t desc_kmajor = make_smem_desc_sw128(0, 16, 1024);
      // V: 16 K2 rows per step = 16 * 128 B inside each 64-column slab; slabs (N blocks) are
      // kChunkBytes apart, 8-row groups 1024 B apart.
      const uint64_t desc_v = make_smem_desc_sw128(0, Cfg::kChunkBytes, 1024);
      // The empty asm makes the bases opaque per call, so the compiler re-derives the
      // per-MMA operands with immediate adds instead of hoisting ~60 loop invariants out of
      // the K2 loop and spilling them (this thread runs with 56 registers: 168*384 allocated at launch = 56*128 + 224*256).
      auto issue_s = [&](int t, uint32_t k_addr) {
        uint32_t tb = tmem_base, qa = smem_q;
        asm volatile("" : "+r"(tb), "+r"(qa), "+r"(k_addr));
        const uint64_t da0 = desc_kmajor + ((qa + t * Cfg::kTileBytes) >> 4);
        const uint64_t db0 = desc_kmajor + (k_addr >> 4);
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk) {
          const uint32_t off = ((kk / 4) * Cfg::kChunkBytes + (kk % 4) * 32) >> 4;
          umma_ss<1, KIND_F16>(tb + t * BN, da0 + off, db0 + off, kIdescS, kk != 0);
        }
      };
      auto issue_pv = [&](int t, uint32_t v_addr, bool accumulate, int half) {
        uint32_t tb = tmem_base;
        asm volatile("" : "+r"(tb), "+r"(v_addr));
        const uint64_t db0 = desc_v + (v_addr >> 4);
#pragma unroll
        for (int kk = half * (BN / 32); kk < (half + 1) * (BN / 32); ++kk) {
          umma_ts_f16<1>(tb + 256 + t * D, tb + t * BN + kk * 8,
                         db0 + (uint32_t)(kk * (2048 >> 4)), kIdescPV,
                         (accumulate || kk != 0) ? 1u : 0u);
        }
      };
      mbar_wait(bar_q_scaled, 0);
      mbar_wait(bar_q_scaled + 8, 0);
      mbar_wait(bar_k_full, 0);
      tcgen05_fence_after();
      issue_s(0, smem_k);
      umma_commit_1cta(bar_s_full);
      issue_s(1, smem_k);
      umma_commit_1cta(bar_s_full + 8);
      umma_commit_1cta(bar_k_empty);
      uint32_t ks = (1 % Cfg::kKStages), kphase = (Cfg::kKStages == 1) ? 1u : 0u;
      uint32_t vs = 0, vphase = 0;
      for (uint32_t j = 0; j < num_kv; ++j) {
        const bool has_next = j + 1 < num_kv;
        const uint32_t v_addr = smem_v + vs * Cfg::kTileBytes;
        const uint32_t k_addr = smem_k + ks * Cfg::kTileBytes;
        mbar_wait(bar_v_full + 8 * vs, vphase);
        mbar_wait(bar_p_full, j & 1);
        tcgen05_fence_after();
        issue_pv(0, v_addr, j != 0, 0);
        mbar_wait(bar_p_full + 8, j & 1);
        tcgen05_fence_after();
        issue_pv(0, v_addr, j != 0, 1);
        if (has_next) {
          mbar_wait(bar_k_full + 8 * ks, kphase);
          tcgen05_fence_after();
          issue_s(0, k_addr);
          umma_commit_1cta(bar_s_full);
        }
        mbar_wait(bar_p_full + 16, j & 1);
        tcgen05_fence_after();
        issue_pv(1, v_addr, j != 0, 0);
        mbar_wait(bar_p_full + 24, j & 1);
        tcgen05_fence_after();
        issue_pv(1, v_addr, j != 0, 1);
        umma_commit_1cta(bar_v_empty + 8 * vs);
        if (has_next) {
          issue_s(1, k_addr);
          umma_commit_1cta(bar_s_full + 8);
          umma_commit_1cta(bar_k_empty + 8 * ks);
          if (++ks == Cfg::kKStages) {
            ks = 0;
            kphase ^= 1;
          }
        }
        if (++vs == Cfg::kVStages) {
          vs = 0;
          vphase ^= 1;
        }
      }
      umma_commit_1cta(bar_o_done);
    }
  } else if (warp >= 4) {
    //===------------------------------------------------------------------===//
    // Softmax warpgroups: tile t = (warp - 4) / 4, one thread per row.
    //===------------------------------------------------------------------===//
    asm volatile("setmaxnreg.inc.sync.aligned.u32 224;");
    const int t = (int)(warp - 4) >> 2;
    const uint32_t quarter = warp & 3;                 // TMEM lane quarter of this warp
    const uint32_t wg_tid = (warp - 4 - 4 * t) * 32 + lane;  // 0..127 within the warpgroup
    const uint32_t row = quarter * 32 + lane;
    const uint32_t lane_base = (quarter * 32u) << 16;
    const uint32_t tmem_s = tmem_base + lane_base + t * BN;
    const uint32_t tmem_o = tmem_base + lane_base + 256 + t * D;

    // ---- Q' = round_T(s' * Q), in place in shared memory (swizzle-agnostic) ----
    {
      const float mult = attention_q_multiplier<kBf16>(__int_as_float(scale_bits));
      mbar_wait(bar_q_full + 8 * t, 0);
      uint8_t* q_tile = smem_gen + (smem_q - smem_base) + t * Cfg::kTileBytes;
#pragma unroll 4
      for (int i = 0; i < Cfg::kTileBytes / 16 / 128; ++i) {
        uint4* p = reinterpret_cast<uint4*>(q_tile) + i * 128 + wg_tid;
        uint4 v = *p;
        uint32_t* w = reinterpret_cast<uint32_t*>(&v);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float lo = elem_to_f32<kBf16>((uint16_t)(w[e] & 0xffffu)) * mult;
          const float hi = elem_to_f32<kBf16>((uint16_t)(w[e] >> 16)) * mult;
          w[e] = pack2<kBf16>(lo, hi);
        }
        *p = v;
      }
      fence_proxy_async_smem();  // generic-proxy writes -> visible to tcgen05.mma
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_q_scaled + 8 * t);
    }

    float m_run = -FLT_MAX;  // TileAttention.cpp:82-84 (useOnlyFiniteValue)
    float l_run = 0.f;
    const int gm_row = m_cta + t * BM + (int)row;
    const uint8_t* mask_row = nullptr;
    if constexpr (kMask) {
      mask_row = mask + ((size_t)batch * M + (size_t)min(gm_row, M - 1)) * (size_t)K2;
    }
    for (uint32_t j = 0; j < num_kv; ++j) {
      const int col0 = (int)((kv_begin + j) * BN);
      // The row's 128 mask bytes of this block, requested before the wait for S.
      uint4 mbytes[kMask ? BN / 16 : 1];
      if constexpr (kMask) {
#pragma unroll
        for (int c = 0; c < BN / 16; ++c) {
          mbytes[c] = make_uint4(0x01010101u, 0x01010101u, 0x01010101u, 0x01010101u);
          if (col0 + c * 16 < K2) {  // K2 % 16 == 0 (selection guard)
            mbytes[c] = ld_nc_na_v4(mask_row + col0 + c * 16);
          }
        }
      }
      mbar_wait(bar_s_full + 8 * t, j & 1);
      tcgen05_fence_after();
      uint32_t s[BN];
#pragma unroll
      for (int c = 0; c < BN / 32; ++c) {
        tmem_ld_32x32b_x32(tmem_s + c * 32, *reinterpret_cast<uint32_t(*)[32]>(&s[c * 32]));
      }
      tmem_wait_ld();
      // Columns >= valid are TMA zero fill (K2 tail) or beyond the row's causal limit.
      int valid = K2 - col0;
      if (causal) valid = min(valid, gm_row + causal_shift + 1 - col0);
      if (valid < BN) {
#pragma unroll
        for (int i = 0; i < BN; ++i) {
          if (i >= valid) s[i] = 0xff800000u;  // -inf
        }
      }
      if constexpr (kMask) {
#pragma unroll
        for (int c = 0; c < BN / 16; ++c) {
          const uint32_t w[4] = {mbytes[c].x, mbytes[c].y, mbytes[c].z, mbytes[c].w};
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            // i1 is stored one byte per element; any non-zero byte is "true".
            if (((w[i >> 2] >> (8 * (i & 3))) & 0xffu) == 0) s[c * 16 + i] = 0xff800000u;
          }
        }
      }
      // Row max: 8 independent chains of 3-input max (a single chain is 128 dependent ops).
      float mxs[8];
#pragma unroll
      for (int a = 0; a < 8; ++a) mxs[a] = max3(__uint_as_float(s[a]), __uint_as_float(s[a + 8]),
                                                __uint_as_float(s[a + 16]));
#pragma unroll
      for (int i = 24; i + 16 <= BN; i += 16) {
#pragma unroll
        for (int a = 0; a < 8; ++a) {
          mxs[a] = max3(mxs[a], __uint_as_float(s[i + a]), __uint_as_float(s[i + a + 8]));
        }
      }
#pragma unroll
      for (int a = 0; a < 8; ++a) mxs[a] = fmaxf(mxs[a], __uint_as_float(s[BN - 8 + a]));
      const float mx = fmaxf(max3(mxs[0], mxs[1], mxs[2]),
                             max3(max3(mxs[3], mxs[4], mxs[5]), mxs[6], mxs[7]));
      const float m_new = fmaxf(m_run, mx);
      // Lazy rescale: keep the stale max while it is within 2^8 of the true one.
      const bool grow = (m_new - m_run) > 8.0f;
      if (j != 0 && __any_sync(0xffffffffu, grow)) {
        const float factor = grow ? fast_exp2(m_run - m_new) : 1.0f;
        l_run *= factor;
#pragma unroll
        for (int c = 0; c < D / 32; ++c) {
          uint32_t o[32];
          tmem_ld_32x32b_x32(tmem_o + c * 32, o);
          tmem_wait_ld();
#pragma unroll
          for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * factor);
          tmem_st_32x32b_x32(tmem_o + c * 32, o);
        }
      }
      if (grow) m_run = m_new;
      // P in two halves of 64 K2 columns: the MMA warp starts P.V on the first half while
      // the second is still on the MUFU (one barrier per half: a waiter may not fall two
      // phases behind on one mbarrier).
      float sum0 = 0.f, sum1 = 0.f;
#pragma unroll
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t p[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const float p0 = fast_exp2(__uint_as_float(s[c * 32 + 2 * i]) - m_run);
          const float x1 = __uint_as_float(s[c * 32 + 2 * i + 1]) - m_run;
          const float p1 = (kPoly == 2 || (kPoly == 1 && (i & 1))) ? poly_exp2(x1) : fast_exp2(x1);
          sum0 += p0;  // f32 P summed before the truncation (:598-601)
          sum1 += p1;
          p[i] = pack2<kBf16>(p0, p1);
        }
        tmem_st_32x32b_x16(tmem_s + c * 16, p);
        if (c & 1) {
          tmem_wait_st();
          tcgen05_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(bar_p_full + 16 * t + 8 * (c >> 1));
        }
      }
      l_run += sum0 + sum1;
    }

    // ---- epilogue: O = cast((1/sum) * acc) ----
    mbar_wait(bar_o_done, 0);
    tcgen05_fence_after();
    if (kSplit && record >= 0) {
      // Split CTA: leave (acc, max, sum) for the merge kernel. m_run may be the stale max of
      // the lazy rescale; acc and l_run are relative to it, so the record is consistent.
      constexpr int kRows = Cfg::kTilesPerCta * BM;
      float* rec = workspace + (size_t)record * (kRows * (D + 2));
      const int r = t * BM + (int)row;
      float4* arow = reinterpret_cast<float4*>(rec + (size_t)r * D);
#pragma unroll
      for (int c = 0; c < D / 32; ++c) {
        uint32_t o[32];
        tmem_ld_32x32b_x32(tmem_o + c * 32, o);
        tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          arow[c * 8 + i] = make_float4(__uint_as_float(o[4 * i]), __uint_as_float(o[4 * i + 1]),
                                        __uint_as_float(o[4 * i + 2]), __uint_as_float(o[4 * i + 3]));
        }
      }
      rec[kRows * D + r] = m_run;
      rec[kRows * (D + 1) + r] = l_run;
    } else {
    const float rcp = 1.0f / l_run;
    const int gm = gm_row;
    uint16_t* orow = reinterpret_cast<uint16_t*>(out) + ((size_t)batch * M + gm) * D;
#pragma unroll
    for (int c = 0; c < D / 32; ++c) {
      uint32_t o[32];
      tmem_ld_32x32b_x32(tmem_o + c * 32, o);
      tmem_wait_ld();
      if (gm < M) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          uint4 v;
          v.x = pack2<kBf16>(__uint_as_float(o[8 * i + 0]) * rcp, __uint_as_float(o[8 * i + 1]) * rcp);
          v.y = pack2<kBf16>(__uint_as_float(o[8 * i + 2]) * rcp, __uint_as_float(o[8 * i + 3]) * rcp);
          v.z = pack2<kBf16>(__uint_as_float(o[8 * i + 4]) * rcp, __uint_as_float(o[8 * i + 5]) * rcp);
          v.w = pack2<kBf16>(__uint_as_float(o[8 * i + 6]) * rcp, __uint_as_float(o[8 * i + 7]) * rcp);
          *reinterpret_cast<uint4*>(orow + c * 32 + i * 8) = v;
        }
      }
    }
    }  // !split
  }

  __syncwarp();
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 2) {
    tcgen05_fence_after();
    tmem_dealloc<1>(tmem_base, Cfg::kTmemCols);
  }
}

// `<kernel>__tma` recipe. Bindings: 0=Q 1=K 2=V 3=O. Push constants: 0=B 1=M 2=K2
// 3=scale (f32 bits) 4=flags.
template <class Cfg, bool kBf16>
constexpr iree_b200_tma_recipe_t make_attention_prefill_recipe() {
  iree_b200_tma_recipe_t r = {};
  r.magic = IREE_B200_TMA_RECIPE_MAGIC;
  r.map_count = 3;
  for (int i = 0; i < 3; ++i) {
    r.maps[i].binding = i;
    r.maps[i].elem_type = kBf16 ? IREE_B200_TMA_BF16 : IREE_B200_TMA_F16;
    r.maps[i].elem_size = 2;
    r.maps[i].rank = 3;
    r.maps[i].dim[0] = IREE_B200_TMA_DIM_LITERAL(Cfg::kHeadDim);
    r.maps[i].dim[1] = IREE_B200_TMA_DIM_CONST(i == 0 ? 1 : 2);
    r.maps[i].dim[2] = IREE_B200_TMA_DIM_CONST(0);
    r.maps[i].box[0] = Cfg::kChunkElems;
    r.maps[i].box[1] = 128;
    r.maps[i].box[2] = 1;
    r.maps[i].swizzle = IREE_B200_TMA_SWIZZLE_128B;
  }
  return r;
}

}  // namespace b200
