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
// One CTA = 256 threads owns a 128 x 256 tile (cta_group::1):
//   warp 0   TMA producer (A 16 KiB + B 32 KiB per stage)
//   warp 1   MMA issuer: per stage 6 tcgen05.cp (2 SFA + 4 SFB atoms) then 4 MMAs
//   warp 2   TMEM allocator         warp 3   scale loader
//   warps 4-7 epilogue (tcgen05.ld -> swizzled smem -> TMA store), as gemm_sm100.cuh
// TMEM: accumulator 256 columns, then per ring stage 8 (SFA) + 16 (SFB) columns. One
// accumulator stage: the scale columns take the room of a second one, so the epilogue of a
// tile is not overlapped with the next main loop (first version).
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

B200_DEVICE void tmem_cp_32x128b_warpx4(uint32_t taddr, uint64_t sdesc) {
  asm volatile("tcgen05.cp.cta_group::1.32x128b.warpx4 [%0], %1;" ::"r"(taddr), "l"(sdesc)
               : "memory");
}

B200_DEVICE void umma_mxf4_ss(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                              uint32_t tmem_sfa, uint32_t tmem_sfb, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::mxf4.block_scale.block32 [%0], %1, %2, %3, [%5], [%6], p;\n\t}" ::
          "r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate), "r"(tmem_sfa), "r"(tmem_sfb)
      : "memory");
}

B200_DEVICE void cp_async_4(uint32_t smem_dst, const void* gmem_src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_dst), "l"(gmem_src) : "memory");
}

B200_DEVICE void cp_async_mbar_arrive_noinc(uint32_t bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
}

using Mx = MxGemmConfig;

__device__ __forceinline__ void mx_gemm_sm100_body(const uint8_t* __restrict__ lhs_scales,
                                                   const uint8_t* __restrict__ rhs_scales,
                                                   void* __restrict__ out, int M, int N, int K,
                                                   int flags, const TensorMap* tm_a,
                                                   const TensorMap* tm_b, const TensorMap* tm_c) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t smem_a = smem_base;
  const uint32_t smem_b = smem_a + Mx::kStages * Mx::kABytes;
  const uint32_t smem_sfa = smem_b + Mx::kStages * Mx::kBBytes;
  const uint32_t smem_sfb = smem_sfa + Mx::kStages * Mx::kSfaBytes;
  const uint32_t smem_epi = smem_base + Mx::kStages * Mx::kStageBytes;
  const uint32_t bar_base = smem_base + Mx::kSmemTiles;
  const uint32_t bar_full = bar_base;                       // [kStages]
  const uint32_t bar_empty = bar_full + 8 * Mx::kStages;    // [kStages]
  const uint32_t bar_tfull = bar_empty + 8 * Mx::kStages;   // [1]
  const uint32_t bar_tempty = bar_tfull + 8;                // [1]
  const uint32_t tmem_slot = bar_tempty + 8;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));

  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t lane = threadIdx.x & 31;
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
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < Mx::kStages; ++s) {
      mbar_init(bar_full + 8 * s, 1 + 32);  // the producer's expect_tx arrive + 32 scale lanes
      mbar_init(bar_empty + 8 * s, 1);      // one tcgen05.commit
    }
    mbar_init(bar_tfull, 1);
    mbar_init(bar_tempty, 4);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<1>(tmem_slot, Mx::kTmemCols);
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_gen + (tmem_slot - smem_base));

  if (warp == 0) {
    //===------------------------------------------------------------------===//
    // TMA producer: operand bytes
    //===------------------------------------------------------------------===//
    if (elect_one()) {
      uint32_t stage = 0, phase = 0;
      for (uint32_t tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        int m_blk, n_blk;
        tile_coord(tile, &m_blk, &n_blk);
        for (uint32_t kb = 0; kb < num_kb; ++kb) {
          mbar_wait(bar_empty + 8 * stage, phase ^ 1);
          mbar_expect_tx(bar_full + 8 * stage, Mx::kABytes + Mx::kBBytes);
          const int k0 = (int)kb * Mx::kBlockKBytes;
          tma_load_3d<1>(smem_a + stage * Mx::kABytes, tm_a, bar_full + 8 * stage, k0,
                         m_blk * Mx::kTileM, 0);
          tma_load_3d<1>(smem_b + stage * Mx::kBBytes, tm_b, bar_full + 8 * stage, k0,
                         n_blk * Mx::kTileN, 0);
          tma_load_3d<1>(smem_b + stage * Mx::kBBytes + Mx::kBBytes / 2, tm_b, bar_full + 8 * stage,
                         k0, n_blk * Mx::kTileN + 128, 0);
          if (++stage == Mx::kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 3) {
    //===------------------------------------------------------------------===//
    // Scale loader: [rows, K/32] row-major -> the instruction's 512-byte atoms
    //===------------------------------------------------------------------===//
    uint32_t stage = 0, phase = 0;
    for (uint32_t tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      int m_blk, n_blk;
      tile_coord(tile, &m_blk, &n_blk);
      for (uint32_t kb = 0; kb < num_kb; ++kb) {
        mbar_wait(bar_empty + 8 * stage, phase ^ 1);
        // 12 rows per lane: rows [0,128) of A then [0,256) of B; two K atoms per row.
#pragma unroll
        for (int i = 0; i < 12; ++i) {
          const int r = i * 32 + (int)lane;         // 0..383
          const bool is_a = r < Mx::kTileM;
          const int local = is_a ? r : r - Mx::kTileM;
          const int global_row = is_a ? min(m_blk * Mx::kTileM + local, M - 1)
                                      : min(n_blk * Mx::kTileN + local, N - 1);
          const uint8_t* src = (is_a ? lhs_scales : rhs_scales) + (size_t)global_row * sf_pitch +
                               (size_t)kb * 8;
          const int in_atom = ((local & 127) & 31) * 16 + ((local & 127) >> 5) * 4;
          // A: atoms [k]; B: atoms [k][n half].
          const uint32_t base = is_a ? smem_sfa + stage * Mx::kSfaBytes
                                     : smem_sfb + stage * Mx::kSfbBytes + (local >> 7) * Mx::kAtomBytes;
          const uint32_t k_stride = is_a ? Mx::kAtomBytes : 2 * Mx::kAtomBytes;
          cp_async_4(base + in_atom, src);
          cp_async_4(base + k_stride + in_atom, src + 4);
        }
        cp_async_mbar_arrive_noinc(bar_full + 8 * stage);
        if (++stage == Mx::kStages) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    //===------------------------------------------------------------------===//
    // MMA issuer
    //===------------------------------------------------------------------===//
    if (elect_one()) {
      uint32_t stage = 0, phase = 0, tphase = 0;
      for (uint32_t tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        mbar_wait(bar_tempty, tphase ^ 1);
        tcgen05_fence_after();
        for (uint32_t kb = 0; kb < num_kb; ++kb) {
          mbar_wait(bar_full + 8 * stage, phase);
          tcgen05_fence_after();
          fence_proxy_async_smem();  // cp.async (generic proxy) scales -> tcgen05.cp (async proxy)
          const uint32_t sfa_col = tmem_base + Mx::kSfCol0 + stage * Mx::kSfColsPerStage;
          const uint32_t sfb_col = sfa_col + 8;
#pragma unroll
          for (int a = 0; a < 2; ++a) {
            tmem_cp_32x128b_warpx4(sfa_col + a * 4,
                                   make_smem_desc_sf_atom(smem_sfa + stage * Mx::kSfaBytes + a * Mx::kAtomBytes));
          }
#pragma unroll
          for (int a = 0; a < 4; ++a) {
            tmem_cp_32x128b_warpx4(sfb_col + a * 4,
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
            umma_mxf4_ss(tmem_base, da, db, idesc, (sfa_col + (j >> 1) * 4) | (sf_id << 30),
                         (sfb_col + (j >> 1) * 8) | (sf_id << 30), (kb | (uint32_t)j) != 0);
          }
          umma_commit_1cta(bar_empty + 8 * stage);
          if (++stage == Mx::kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit_1cta(bar_tfull);
        tphase ^= 1;
      }
    }
  } else if (warp >= 4) {
    //===------------------------------------------------------------------===//
    // Epilogue
    //===------------------------------------------------------------------===//
    const uint32_t ew = warp - 4;
    const uint32_t row = ew * 32 + lane;
    const bool epi_leader = (warp == 4) && (lane == 0);
    const bool accumulate = (flags & 1) != 0;
    uint32_t tphase = 0, buf = 0;
    for (uint32_t tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      int m_blk, n_blk;
      tile_coord(tile, &m_blk, &n_blk);
      const int m0 = m_blk * Mx::kTileM, n0 = n_blk * Mx::kTileN;
      mbar_wait(bar_tfull, tphase);
      tcgen05_fence_after();
      const uint32_t taddr0 = tmem_base + ((ew * 32u) << 16);
#pragma unroll 1
      for (int c = 0; c < Mx::kTileN / Mx::kEpiCols; ++c) {
        if (epi_leader) tma_store_wait_read<Mx::kEpiBufs - 1>();
        named_barrier_sync(1, 128);
        uint32_t v[32];
        tmem_ld_32x32b_x32(taddr0 + c * Mx::kEpiCols, v);
        tmem_wait_ld();
        if (c == Mx::kTileN / Mx::kEpiCols - 1) {
          tcgen05_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(bar_tempty);
        }
        if (accumulate) {
          const int gm = m0 + (int)row, gn = n0 + c * Mx::kEpiCols;
          if (gm < M) {
            const float* cin = reinterpret_cast<const float*>(out) + (size_t)gm * (size_t)N + gn;
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (gn + i < N) v[i] = __float_as_uint(__uint_as_float(v[i]) + cin[i]);
          }
        }
        const uint32_t row_addr = smem_epi + buf * Mx::kEpiBytes + row * 128;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const uint32_t dst = row_addr + ((i ^ (row & 7)) << 4);
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "r"(v[4 * i]),
                       "r"(v[4 * i + 1]), "r"(v[4 * i + 2]), "r"(v[4 * i + 3])
                       : "memory");
        }
        fence_proxy_async_smem();
        named_barrier_sync(1, 128);
        if (epi_leader) {
          tma_store_3d(tm_c, smem_epi + buf * Mx::kEpiBytes, n0 + c * Mx::kEpiCols, m0, 0);
          tma_store_commit();
        }
        buf ^= 1;
      }
      tphase ^= 1;
    }
    if (epi_leader) tma_store_wait<0>();
  }

  __syncwarp();
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 2) {
    tcgen05_fence_after();
    tmem_dealloc<1>(tmem_base, Mx::kTmemCols);
  }
}

// `iree_b200_matmul_mxfp4_f32__tma`. Bindings: 0 lhs, 1 lhs_scales, 2 rhs, 3 rhs_scales,
// 4 out. Push constants: 0 M, 1 N, 2 K, 3 flags, 4 K / 2 (packed bytes per row).
constexpr iree_b200_tma_recipe_t make_mx_gemm_recipe() {
  iree_b200_tma_recipe_t r = {};
  r.magic = IREE_B200_TMA_RECIPE_MAGIC;
  r.map_count = 3;
  for (int i = 0; i < 2; ++i) {  // packed operand bytes [rows, K/2]
    r.maps[i].binding = i == 0 ? 0 : 2;
    r.maps[i].elem_type = IREE_B200_TMA_U8;
    r.maps[i].elem_size = 1;
    r.maps[i].rank = 3;
    r.maps[i].dim[0] = IREE_B200_TMA_DIM_CONST(4);
    r.maps[i].dim[1] = IREE_B200_TMA_DIM_CONST(i == 0 ? 0 : 1);
    r.maps[i].dim[2] = IREE_B200_TMA_DIM_LITERAL(1);
    r.maps[i].box[0] = MxGemmConfig::kBlockKBytes;
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
  r.maps[2].box[0] = MxGemmConfig::kEpiCols;
  r.maps[2].box[1] = MxGemmConfig::kTileM;
  r.maps[2].box[2] = 1;
  r.maps[2].swizzle = IREE_B200_TMA_SWIZZLE_128B;
  return r;
}

}  // namespace b200
