// mx_gemm_sm100.cuh — MXFP4 matmul on the block-scaled tensor-core path of sm_100a:
// tcgen05.mma.kind::mxf4.block_scale (f4E2M1FN operands, one f8E8M0FNU scale per 32 elements,
// f32 accumulate in TMEM), scales moved shared memory -> TMEM with tcgen05.cp.
//
// Semantics: tests/e2e/matmul/generate_code_mx.py:153-172 (oracle.c oracle_matmul_mxfp4_f32)
//   out[m, n] = (out[m, n] +) sum_kb sum_i (f4(lhs[m,kb,i]) * 2^(ls[m,kb]-127)) *
//                                          (f4(rhs[n,kb,i]) * 2^(rs[n,kb]-127))
// lhs [M, K/2] and rhs [N, K/2] hold two f4E2M1FN per byte, element 2i in the low nibble
// (mmt4d_tile_generic.c:38-45) — which is the packed K-major operand layout the instruction
// reads from shared memory, so the operand tiles are moved as bytes by TMA (128 bytes = 256
// elements of K per row and stage, SWIZZLE_128B, exactly the geometry of the i8 kernel).
// Products of two scaled f4 values are exact in f32 and the accumulator is f32: the result
// equals the reference for every input whose partial sums are exact and differs by f32
// summation order otherwise, like every tcgen05 GEMM here.
//
// Scale factors. The instruction (block32: K = 64 per MMA, 2 scales per row) reads them from
// TMEM in a fixed layout: the scales of 128 rows x 4 consecutive K blocks form one 512-byte
// atom, byte offset (r % 32) * 16 + (r / 32) * 4 + c in shared memory; one
// tcgen05.cp.32x128b.warpx4 turns an atom into 4 TMEM columns (lane = r % 32, column =
// r / 32, byte = c, replicated to the four lane quarters). MMA j of a stage takes the atom
// j / 2 and, through the scale-factor id in the instruction descriptor (bits [4,6) for B,
// [29,31) for A — the byte offset inside the 32-bit column, also carried in the two top
// bits of the TMEM address operand), bytes 2 * (j % 2) .. + 1. The reference keeps scales
// row-major [rows, K/32], so a loader warp gathers them: per row and stage 8 bytes, as two
// 4-byte cp.async into the two K atoms, completion tracked by the stage's full barrier
// (cp.async.mbarrier.arrive.noinc) together with the TMA bytes.
//
// One CTA = 384 threads; cta_group::2 pairs two CTAs of a cluster on a 256 x 256 tile (each
// holds 128 rows of A, 128 of the 256 columns of B, and ALL 256 columns' scales: both tensor
// cores need them), cta_group::1 owns 128 x 256 (small grids):
//   warp 0    TMA producer (A 16 KiB + B 16 / 32 KiB per stage)
//   warp 1    MMA issuer (pair leader): per stage 6 tcgen05.cp (2 SFA + 4 SFB atoms), 4 MMAs
//   warp 2    TMEM allocator
//   warp 3    scale loader: cp.async gather kScaleLookahead stages ahead; when a stage's
//             scales have landed (its own mbarrier) it fences them to the async proxy and
//             arrives on the LEADER's full barrier (remote arrive from the peer CTA)
//   warps 4-11 epilogue, two warpgroups (column halves 0-127 / 128-255): each thread pulls
//             its row's 128 accumulator columns into registers at once and hands TMEM back
//             (~0.5 k cycles), then stages 32-column chunks through shared memory to TMA
//             stores while the next tile's main loop is already running — the single
//             accumulator stage (the scale columns take the room of a second one) costs no
//             overlap.
// TMEM: accumulator 256 columns, then per ring stage 8 (SFA) + 16 (SFB) columns.
// Guards (ops layer): K % 256 == 0, 16-byte aligned operand rows, N % 4 == 0.

#pragma once

#include "common.cuh"
#include "gemm_config.h"
#include "iree_b200_tma_recipe.h"

namespace b200 {

// No-swizzle shared-memory matrix descriptor of one scale atom: 32 rows of 16 bytes, 8-row
// core matrices 128 bytes apart.
B200_DEVICE uint64_t make_smem_desc_sf_atom(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>(16 >> 4) << 16;   // leading byte offset (unused: one 16-byte column)
  d |= static_cast<uint64_t>(128 >> 4) << 32;  // stride byte offset between 8-row groups
  d |= 1ull << 46;
  return d;
}

template <int kCtaGroup>
B200_DEVICE void tmem_cp_32x128b_warpx4(uint32_t taddr, uint64_t sdesc) {
  if constexpr (kCtaGroup == 1) {
    asm volatile("tcgen05.cp.cta_group::1.32x128b.warpx4 [%0], %1;" ::"r"(taddr), "l"(sdesc)
                 : "memory");
  } else {
    asm volatile("tcgen05.cp.cta_group::2.32x128b.warpx4 [%0], %1;" ::"r"(taddr), "l"(sdesc)
                 : "memory");
  }
}

template <int kCtaGroup>
B200_DEVICE void umma_mxf4_ss(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                              uint32_t tmem_sfa, uint32_t tmem_sfb, uint32_t accumulate) {
  if constexpr (kCtaGroup == 1) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::mxf4.block_scale.block32 [%0], %1, %2, %3, [%5], [%6], p;\n\t}" ::
            "r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate), "r"(tmem_sfa), "r"(tmem_sfb)
        : "memory");
  } else {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::mxf4.block_scale.block32 [%0], %1, %2, %3, [%5], [%6], p;\n\t}" ::
            "r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate), "r"(tmem_sfa), "r"(tmem_sfb)
        : "memory");
  }
}

B200_DEVICE void cp_async_4(uint32_t smem_dst, const void* gmem_src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_dst), "l"(gmem_src) : "memory");
}

B200_DEVICE void cp_async_mbar_arrive_noinc(uint32_t bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
}

// kPacked: the scale bindings already hold the instruction's atoms, [ceil(rows / 128)]
// [K / 128][512 bytes] (iree_b200_mxfp4_pack_scales, or a module whose encoding packed them at
// compile time): the TMA producer brings them in with two more tensor maps and the loader
// warp has nothing to do. Otherwise they are row-major [rows, K / 32] and gathered in the
// kernel — no workspace, but the 4-byte gathers bound the kernel at about half the rate.
template <class Mx, bool kPacked>
__device__ __forceinline__ void mx_gemm_sm100_body(const uint8_t* __restrict__ lhs_scales,
                                                   const uint8_t* __restrict__ rhs_scales,
                                                   void* __restrict__ out, int M, int N, int K,
                                                   int flags, const TensorMap* tm_a,
                                                   const TensorMap* tm_b, const TensorMap* tm_c,
                                                   const TensorMap* tm_sfa = nullptr,
                                                   const TensorMap* tm_sfb = nullptr) {
  constexpr int G = Mx::kCtaGroup;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t smem_a = smem_base;
  const uint32_t smem_b = smem_a + Mx::kStages * Mx::kABytes;
  const uint32_t smem_sfa = smem_b + Mx::kStages * Mx::kBBytes;
  const uint32_t smem_sfb = smem_sfa + Mx::kStages * Mx::kSfaBytes;
  const uint32_t smem_epi = smem_base + Mx::kStages * Mx::kStageBytes;
  const uint32_t bar_base = smem_base + Mx::kSmemTiles;
  const uint32_t bar_full = bar_base;                        // [kStages] (the leader's are used)
  const uint32_t bar_empty = bar_full + 8 * Mx::kStages;     // [kStages]
  const uint32_t bar_sf = bar_empty + 8 * Mx::kStages;       // [kStages] this CTA's scales landed
  const uint32_t bar_tfull = bar_sf + 8 * Mx::kStages;       // [1]
  const uint32_t bar_tempty = bar_tfull + 8;                 // [1]
  const uint32_t tmem_slot = bar_tempty + 8;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));

  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t lane = threadIdx.x & 31;
  const uint32_t cta_rank = (G == 2) ? cluster_ctarank() : 0u;
  const bool is_leader = cta_rank == 0;
  const uint32_t cluster_idx = (G == 2) ? cluster_id_x() : blockIdx.x;
  const uint32_t cluster_cnt = (G == 2) ? num_clusters_x() : gridDim.x;
  const uint32_t tiles_m = (M + Mx::kTileM - 1) / Mx::kTileM;
  const uint32_t tiles_n = (N + Mx::kTileN - 1) / Mx::kTileN;
  const uint32_t total_tiles = tiles_m * tiles_n;
  const uint32_t num_kb = (uint32_t)K / Mx::kBlockK;   // K % 256 == 0 (guard)
  const int sf_pitch = K / 32;                          // scale bytes per row

  auto tile_coord = [&](uint32_t tile, int* m_blk, int* n_blk) {
    const uint32_t group_span = Mx::kGroupM * tiles_n;
    const uint32_t group = tile / group_span;
    const uint32_t first_m = group * Mx::kGroupM;
    const uint32_t group_m = min(tiles_m - first_m, (uint32_t)Mx::kGroupM);
    const uint32_t in_group = tile % group_span;
    *m_blk = (int)(first_m + in_group % group_m);
    *n_blk = (int)(in_group / group_m);
  };

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(tm_a);
    tma_prefetch_desc(tm_b);
    tma_prefetch_desc(tm_c);
    if constexpr (kPacked) {
      tma_prefetch_desc(tm_sfa);
      tma_prefetch_desc(tm_sfb);
    }
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < Mx::kStages; ++s) {
      // The leader producer's expect_tx (covering both CTAs' TMA bytes) + (gather form) one
      // "scales landed and fenced" arrive per CTA of the pair.
      mbar_init(bar_full + 8 * s, kPacked ? 1 : 1 + G);
      mbar_init(bar_empty + 8 * s, 1);   // one tcgen05.commit (multicast to the pair)
      mbar_init(bar_sf + 8 * s, 32);     // the 32 lanes of this CTA's scale loader
    }
    mbar_init(bar_tfull, 1);
    mbar_init(bar_tempty, 8 * G);        // one arrive per epilogue warp of the pair
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<G>(tmem_slot, Mx::kTmemCols);
  tcgen05_fence_before();
  if constexpr (G == 2) {
    cluster_sync();
  } else {
    __syncthreads();
  }
  tcgen05_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_gen + (tmem_slot - smem_base));

  if (warp < 4) asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");
  if (warp == 0) {
    //===------------------------------------------------------------------===//
    // TMA producer: operand bytes
    //===------------------------------------------------------------------===//
    if (elect_one()) {
      const uint32_t full0 = (G == 2) ? map_to_cta(bar_full, 0) : bar_full;
      uint32_t stage = 0, phase = 0;
      for (uint32_t tile = cluster_idx; tile < total_tiles; tile += cluster_cnt) {
        int m_blk, n_blk;
        tile_coord(tile, &m_blk, &n_blk);
        const int m0 = m_blk * Mx::kTileM + (int)cta_rank * Mx::kBlockM;
        const int n0 = n_blk * Mx::kTileN + (int)cta_rank * Mx::kBlockN;
        for (uint32_t kb = 0; kb < num_kb; ++kb) {
          mbar_wait(bar_empty + 8 * stage, phase ^ 1);
          const uint32_t full = full0 + 8 * stage;
          if (is_leader) {
            mbar_expect_tx(bar_full + 8 * stage,
                           (Mx::kABytes + Mx::kBBytes + (kPacked ? Mx::kSfaBytes + Mx::kSfbBytes : 0)) * G);
          }
          const int k0 = (int)kb * Mx::kBlockKBytes;
          if constexpr (kPacked) {
            // Atoms [row tile][K atom][128 words]: two K atoms of this CTA's A rows, two of each
            // 128-column half of the tile's B columns.
            tma_load_3d<G>(smem_sfa + stage * Mx::kSfaBytes, tm_sfa, full, 0, (int)kb * 2,
                           m_blk * G + (int)cta_rank);
#pragma unroll
            for (int nh = 0; nh < 2; ++nh) {
              tma_load_3d<G>(smem_sfb + stage * Mx::kSfbBytes + nh * 2 * Mx::kAtomBytes, tm_sfb, full, 0,
                             (int)kb * 2, n_blk * 2 + nh);
            }
          }
          tma_load_3d<G>(smem_a + stage * Mx::kABytes, tm_a, full, k0, m0, 0);
#pragma unroll
          for (int j = 0; j < Mx::kBlockN / 128; ++j) {
            tma_load_3d<G>(smem_b + stage * Mx::kBBytes + j * (128 * Mx::kBlockKBytes), tm_b, full,
                           k0, n0 + j * 128, 0);
          }
          if (++stage == Mx::kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 3 && !kPacked) {
    //===------------------------------------------------------------------===//
    // Scale loader: [rows, K/32] row-major -> the instruction's 512-byte atoms
    //===------------------------------------------------------------------===//
    const uint32_t my_tiles =
        cluster_idx < total_tiles ? (total_tiles - cluster_idx + cluster_cnt - 1) / cluster_cnt : 0;
    const uint32_t total_its = my_tiles * num_kb;
    const uint32_t full0 = (G == 2) ? map_to_cta(bar_full, 0) : bar_full;
    int m_blk = 0, n_blk = 0;
    for (uint32_t it = 0; it < total_its + Mx::kScaleLookahead; ++it) {
      if (it < total_its) {
        const uint32_t stage = it % Mx::kStages, phase = (it / Mx::kStages) & 1;
        const uint32_t kb = it % num_kb;
        if (kb == 0) tile_coord(cluster_idx + (it / num_kb) * cluster_cnt, &m_blk, &n_blk);
        mbar_wait(bar_empty + 8 * stage, phase ^ 1);
        // 12 rows per lane: this CTA's 128 rows of A, then all 256 rows of B; two K atoms each.
        // (flags bit 4: diagnostic — no scale loads, results are garbage; tools only.)
#pragma unroll
        for (int i = 0; i < ((flags & 0x10) ? 0 : 12); ++i) {
          const int r = i * 32 + (int)lane;         // 0..383
          const bool is_a = r < Mx::kBlockM;
          const int local = is_a ? r : r - Mx::kBlockM;
          const int global_row =
              is_a ? min(m_blk * Mx::kTileM + (int)cta_rank * Mx::kBlockM + local, M - 1)
                   : min(n_blk * Mx::kTileN + local, N - 1);
          const uint8_t* src = (is_a ? lhs_scales : rhs_scales) + (size_t)global_row * sf_pitch +
                               (size_t)kb * 8;
          const int in_atom = ((local & 127) & 31) * 16 + ((local & 127) >> 5) * 4;
          // A: atoms [k]; B: atoms [n half][k].
          const uint32_t base = is_a ? smem_sfa + stage * Mx::kSfaBytes
                                     : smem_sfb + stage * Mx::kSfbBytes + (local >> 7) * 2 * Mx::kAtomBytes;
          cp_async_4(base + in_atom, src);
          cp_async_4(base + Mx::kAtomBytes + in_atom, src + 4);
        }
        cp_async_mbar_arrive_noinc(bar_sf + 8 * stage);
      }
      if (it >= (uint32_t)Mx::kScaleLookahead) {
        const uint32_t j = it - Mx::kScaleLookahead;
        const uint32_t stage = j % Mx::kStages, phase = (j / Mx::kStages) & 1;
        mbar_wait(bar_sf + 8 * stage, phase);
        fence_proxy_async_smem();  // cp.async (generic proxy) scales -> tcgen05.cp (async proxy)
        __syncwarp();
        if (lane == 0) {
          if constexpr (G == 2) {
            mbar_arrive_cluster(full0 + 8 * stage);
          } else {
            mbar_arrive(bar_full + 8 * stage);
          }
        }
      }
    }
  } else if (warp == 1) {
    //===------------------------------------------------------------------===//
    // MMA issuer (pair leader only)
    //===------------------------------------------------------------------===//
    if (is_leader && elect_one()) {
      uint32_t stage = 0, phase = 0, tphase = 0;
      for (uint32_t tile = cluster_idx; tile < total_tiles; tile += cluster_cnt) {
        mbar_wait(bar_tempty, tphase ^ 1);
        tcgen05_fence_after();
        for (uint32_t kb = 0; kb < num_kb; ++kb) {
          mbar_wait(bar_full + 8 * stage, phase);
          tcgen05_fence_after();
          const uint32_t sfa_col = tmem_base + Mx::kSfCol0 + stage * Mx::kSfColsPerStage;
          const uint32_t sfb_col = sfa_col + 8;
#pragma unroll
          for (int a = 0; a < 2; ++a) {
            tmem_cp_32x128b_warpx4<G>(
                sfa_col + a * 4,
                make_smem_desc_sf_atom(smem_sfa + stage * Mx::kSfaBytes + a * Mx::kAtomBytes));
          }
#pragma unroll
          for (int a = 0; a < 4; ++a) {
            // shared memory order [n half][K atom]; TMEM order [K atom][n half].
            tmem_cp_32x128b_warpx4<G>(
                sfb_col + (a & 1) * 8 + (a >> 1) * 4,
                make_smem_desc_sf_atom(smem_sfb + stage * Mx::kSfbBytes + a * Mx::kAtomBytes));
          }
          const uint32_t a_addr = smem_a + stage * Mx::kABytes;
          const uint32_t b_addr = smem_b + stage * Mx::kBBytes;
#pragma unroll
          for (int j = 0; j < Mx::kKSteps; ++j) {
            const uint64_t da = make_smem_desc_sw128(a_addr + j * 32, 16, 1024);
            const uint64_t db = make_smem_desc_sw128(b_addr + j * 32, 16, 1024);
            const uint32_t sf_id = (uint32_t)(j & 1) * 2u;  // bytes 0-1 or 2-3 of the scale column
            const uint32_t idesc = Mx::kIdesc | (sf_id << 4) | (sf_id << 29);
            umma_mxf4_ss<G>(tmem_base, da, db, idesc, (sfa_col + (j >> 1) * 4) | (sf_id << 30),
                            (sfb_col + (j >> 1) * 8) | (sf_id << 30), (kb | (uint32_t)j) != 0);
          }
          umma_commit<G>(bar_empty + 8 * stage);
          if (++stage == Mx::kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit<G>(bar_tfull);
        tphase ^= 1;
      }
    }
  } else if (warp >= 4) {
    //===------------------------------------------------------------------===//
    // Epilogue: warpgroup h = columns [128 h, 128 h + 128) of the tile
    //===------------------------------------------------------------------===//
    asm volatile("setmaxnreg.inc.sync.aligned.u32 224;");
    const uint32_t h = (warp - 4) >> 2;
    const uint32_t ew = warp & 3;            // TMEM lane quarter of this warp
    const uint32_t row = ew * 32 + lane;     // row within this CTA's 128
    const bool epi_leader = (ew == 0) && (lane == 0);
    const bool accumulate = (flags & 1) != 0;
    const uint32_t tempty0 = (G == 2) ? map_to_cta(bar_tempty, 0) : bar_tempty;
    const uint32_t stage_buf = smem_epi + h * Mx::kEpiBytes;
    uint32_t tphase = 0;
    for (uint32_t tile = cluster_idx; tile < total_tiles; tile += cluster_cnt) {
      int m_blk, n_blk;
      tile_coord(tile, &m_blk, &n_blk);
      const int m0 = m_blk * Mx::kTileM + (int)cta_rank * Mx::kBlockM;
      const int n0 = n_blk * Mx::kTileN + (int)h * 128;
      mbar_wait(bar_tfull, tphase);
      tcgen05_fence_after();
      const uint32_t taddr0 = tmem_base + ((ew * 32u) << 16) + h * 128;
      uint32_t v[128];
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        tmem_ld_32x32b_x32(taddr0 + c * 32, *reinterpret_cast<uint32_t(*)[32]>(&v[c * 32]));
      }
      tmem_wait_ld();
      // The accumulator is in registers: the next tile's MMAs may overwrite TMEM.
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) {
        if constexpr (G == 2) {
          mbar_arrive_cluster(tempty0);
        } else {
          mbar_arrive(bar_tempty);
        }
      }
      if (accumulate) {
        const int gm = m0 + (int)row;
        if (gm < M) {
          const float* cin = reinterpret_cast<const float*>(out) + (size_t)gm * (size_t)N + n0;
#pragma unroll
          for (int i = 0; i < 128; ++i)
            if (n0 + i < N) v[i] = __float_as_uint(__uint_as_float(v[i]) + cin[i]);
        }
      }
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        // The staging buffer was last read by the TMA store of the previous chunk.
        if (epi_leader) tma_store_wait_read<0>();
        named_barrier_sync(1 + h, 128);
        const uint32_t row_addr = stage_buf + row * 128;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const uint32_t dst = row_addr + ((i ^ (row & 7)) << 4);
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "r"(v[c * 32 + 4 * i]),
                       "r"(v[c * 32 + 4 * i + 1]), "r"(v[c * 32 + 4 * i + 2]),
                       "r"(v[c * 32 + 4 * i + 3])
                       : "memory");
        }
        fence_proxy_async_smem();
        named_barrier_sync(1 + h, 128);
        if (epi_leader) {
          tma_store_3d(tm_c, stage_buf, n0 + c * Mx::kEpiCols, m0, 0);
          tma_store_commit();
        }
      }
      tphase ^= 1;
    }
    if (epi_leader) tma_store_wait<0>();
  }

  __syncwarp();
  tcgen05_fence_before();
  if constexpr (G == 2) {
    cluster_sync();
  } else {
    __syncthreads();
  }
  if (warp == 2) {
    tcgen05_fence_after();
    tmem_dealloc<G>(tmem_base, Mx::kTmemCols);
  }
}

// `iree_b200_matmul_mxfp4_f32[_1cta]__tma`. Bindings: 0 lhs, 1 lhs_scales, 2 rhs,
// 3 rhs_scales, 4 out. Push constants: 0 M, 1 N, 2 K, 3 flags, 4 K / 2 (packed bytes per row).
// Packed form: + 5 K / 128, 6 ceil(M / 128), 7 ceil(N / 128), and two more maps over the scale
// atoms viewed as [row tiles][K atoms][128 u32 words].
constexpr iree_b200_tma_recipe_t make_mx_gemm_recipe(bool packed) {
  iree_b200_tma_recipe_t r = {};
  r.magic = IREE_B200_TMA_RECIPE_MAGIC;
  r.map_count = packed ? 5 : 3;
  if (packed) {
    for (int i = 3; i < 5; ++i) {
      r.maps[i].binding = i == 3 ? 1 : 3;
      r.maps[i].elem_type = IREE_B200_TMA_U32;
      r.maps[i].elem_size = 4;
      r.maps[i].rank = 3;
      r.maps[i].dim[0] = IREE_B200_TMA_DIM_LITERAL(128);
      r.maps[i].dim[1] = IREE_B200_TMA_DIM_CONST(5);
      r.maps[i].dim[2] = IREE_B200_TMA_DIM_CONST(i == 3 ? 6 : 7);
      r.maps[i].box[0] = 128;
      r.maps[i].box[1] = 2;
      r.maps[i].box[2] = 1;
      r.maps[i].swizzle = IREE_B200_TMA_SWIZZLE_NONE;
    }
  }
  for (int i = 0; i < 2; ++i) {  // packed operand bytes [rows, K/2]
    r.maps[i].binding = i == 0 ? 0 : 2;
    r.maps[i].elem_type = IREE_B200_TMA_U8;
    r.maps[i].elem_size = 1;
    r.maps[i].rank = 3;
    r.maps[i].dim[0] = IREE_B200_TMA_DIM_CONST(4);
    r.maps[i].dim[1] = IREE_B200_TMA_DIM_CONST(i == 0 ? 0 : 1);
    r.maps[i].dim[2] = IREE_B200_TMA_DIM_LITERAL(1);
    r.maps[i].box[0] = 128;   // kBlockKBytes
    r.maps[i].box[1] = 128;
    r.maps[i].box[2] = 1;
    r.maps[i].swizzle = IREE_B200_TMA_SWIZZLE_128B;
  }
  r.maps[2].binding = 4;
  r.maps[2].elem_type = IREE_B200_TMA_F32;
  r.maps[2].elem_size = 4;
  r.maps[2].rank = 3;
  r.maps[2].dim[0] = IREE_B200_TMA_DIM_CONST(1);
  r.maps[2].dim[1] = IREE_B200_TMA_DIM_CONST(0);
  r.maps[2].dim[2] = IREE_B200_TMA_DIM_LITERAL(1);
  r.maps[2].box[0] = 32;      // kEpiCols
  r.maps[2].box[1] = 128;     // kBlockM
  r.maps[2].box[2] = 1;
  r.maps[2].swizzle = IREE_B200_TMA_SWIZZLE_128B;
  return r;
}

}  // namespace b200
