// gemm_sm100.cuh — persistent, warp-specialised tcgen05 GEMM for sm_100a.
//
// Replaces, for matching dispatches, the kernels IREE's LLVMGPU codegen emits for
// linalg.matmul / linalg.batch_matmul (mma.sync.m16n8k16 schedules chosen in
// compiler/src/iree/compiler/Codegen/LLVMGPU/KernelConfig.cpp:2398-2470). Semantics are
// the reference's own statement of the op
// (tools/testing/e2e/iree-e2e-matmul-test.cc:28-49,73-86,108-121):
//     C[b,m,n] = (flags&1 ? C[b,m,n] : 0) + sum_k ext(A[b,m,k]) * ext(B[b,k,n])
// with f32 accumulation for f16/bf16 and wrapping i32 accumulation for i8.
//
// Layouts (dense row-major, as IREE passes them): A [batch,M,K] (K-major),
// B [batch,K,N] (N contiguous => "MN-major" UMMA operand) or, with kBMajorK,
// B [batch,N,K] (the transpose_rhs form, tests/e2e/matmul/generate_code.py:172-174),
// C [batch,M,N] f32 / i32.
//
// Structure (one CTA = 256 threads; kCtaGroup==2 pairs two CTAs of a cluster on one
// 256x256 tile, each CTA holding 128 accumulator rows in its own TMEM):
//   warp 0      TMA producer: A tile (128 x 128 B, SWIZZLE_128B) and this CTA's share of
//               the B tile into a kStages-deep shared-memory ring, mbarrier-tracked.
//   warp 1      MMA issuer (one elected thread, pair leader only): tcgen05.mma
//               SS-form, accumulators in TMEM, 2 accumulator stages of 256 columns so the
//               epilogue of tile i overlaps the main loop of tile i+1.
//   warp 2      TMEM allocator / deallocator.
//   warps 4..7  epilogue: tcgen05.ld 32 columns at a time -> swizzled smem staging ->
//               TMA store (ragged M/N edges are clipped by the TMA unit; ragged K is
//               zero-filled by it), double-buffered.
// Tiles are walked persistently (grid = #SMs) in an L2-friendly grouped raster.
//
// Roofline: tensor-pipe bound. Algorithmic work 2*batch*M*N*K FLOP; algorithmic bytes
// batch*(M*K*sa + K*N*sb + M*N*4).

#pragma once

#include "common.cuh"
#include "gemm_config.h"
#include "iree_b200_tma_recipe.h"

namespace b200 {

// Two f32 -> one register of two 16-bit elements (lo = first), round to nearest even.
template <bool kBf16>
B200_DEVICE uint32_t pack_pair(float lo, float hi) {
  uint32_t r;
  if constexpr (kBf16) {
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  } else {
    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  }
  return r;
}

struct TileCoord {
  int batch, m_blk, n_blk;
};

template <class Cfg>
B200_DEVICE TileCoord decode_tile(uint32_t tile, uint32_t tiles_m, uint32_t tiles_n,
                                  uint32_t group_rows = Cfg::kGroupM) {
  const uint32_t per_batch = tiles_m * tiles_n;
  TileCoord c;
  c.batch = tile / per_batch;
  const uint32_t t = tile % per_batch;
  const uint32_t group_span = group_rows * tiles_n;
  const uint32_t group = t / group_span;
  const uint32_t first_m = group * group_rows;
  const uint32_t group_m = min(tiles_m - first_m, group_rows);
  const uint32_t in_group = t % group_span;
  c.m_blk = first_m + in_group % group_m;
  c.n_blk = in_group / group_m;
  return c;
}

template <class Cfg, bool kDeviceMaps = false>
__device__ __forceinline__ void gemm_sm100_body(const void* __restrict__ lhs,
                                                const void* __restrict__ rhs, void* __restrict__ out,
                                                int M, int N, int K, int batch, int flags,
                                                const TensorMap* tm_a, const TensorMap* tm_b,
                                                const TensorMap* tm_c,
                                                const void* __restrict__ bias = nullptr,
                                                int tile_m0 = 1, int tile_n0 = 1,
                                                uint32_t epi_swizzle = 7) {
  // kTiled (linalg.mmt4d): M = M1 * M0, N = N1 * N0, K = K1 * K0 are the logical extents,
  // |tile_m0| / |tile_n0| the tile rows of lhs / rhs (= result tile shape), multiples of 8;
  // K0 rows are 16 bytes. 8 tile rows x 16 bytes are contiguous in memory = one UMMA core
  // matrix, so the operands sit in shared memory UNSWIZZLED as [8-row group][8 K tiles]
  // [8 rows][16 B] (make_gemm_tiled_recipe). |epi_swizzle| = the XOR mask of the result
  // map's swizzle (7 / 3 / 1 for 128 / 64 / 32-byte result-tile rows).
  constexpr int G = Cfg::kCtaGroup;
  static_assert(!Cfg::kTiled || (Cfg::kBMajorK && !Cfg::kOut16 && !Cfg::kBias),
                "the tiled form is the NT f32 / i32 GEMM");
  (void)lhs;
  (void)rhs;

  extern __shared__ uint8_t smem_raw[];
  // SWIZZLE_128B tiles need 1024-byte alignment.
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t smem_a = smem_base;
  const uint32_t smem_b = smem_base + Cfg::kStages * Cfg::kABytes;
  const uint32_t smem_epi = smem_base + Cfg::kStages * Cfg::kStageBytes;
  const uint32_t bar_base = smem_base + Cfg::kSmemTiles;
  const uint32_t bar_full = bar_base;                          // [kStages]
  const uint32_t bar_empty = bar_full + 8 * Cfg::kStages;      // [kStages]
  const uint32_t bar_tfull = bar_empty + 8 * Cfg::kStages;     // [kAccStages]
  const uint32_t bar_tempty = bar_tfull + 8 * Cfg::kAccStages; // [kAccStages]
  const uint32_t tmem_slot = bar_tempty + 8 * Cfg::kAccStages; // uint32
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));

  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t lane = threadIdx.x & 31;
  const uint32_t cta_rank = (G == 2) ? cluster_ctarank() : 0u;
  const bool is_leader = cta_rank == 0;
  const uint32_t cluster_idx = (G == 2) ? cluster_id_x() : blockIdx.x;
  const uint32_t cluster_cnt = (G == 2) ? num_clusters_x() : gridDim.x;

  const uint32_t tiles_m = (M + Cfg::kTileM - 1) / Cfg::kTileM;
  const uint32_t tiles_n = (N + Cfg::kTileN - 1) / Cfg::kTileN;
  const uint32_t total_tiles = tiles_m * tiles_n * (uint32_t)batch;
  const uint32_t num_kb = (K + Cfg::kBlockK - 1) / Cfg::kBlockK;
  // Diagnostic bits (tools/gemm_diag.py; never set by the ops layer): results are
  // garbage, the pipeline structure is unchanged.
  const bool diag_no_store = (flags & 0x10) != 0;  // epilogue drains TMEM but stores nothing
  const bool diag_no_load = (flags & 0x20) != 0;   // no TMA loads; MMA does not wait for data
  const bool diag_no_mma = (flags & 0x40) != 0;    // MMA warp only recycles the ring
  const bool diag_l2_hint = (flags & 0x80) != 0;   // A loads evict_last, B loads evict_first
  // bits 8-15: raster group height override (tile rows walked before the column advances)
  const uint32_t group_rows = ((flags >> 8) & 0xff) ? ((flags >> 8) & 0xff) : Cfg::kGroupM;

  // kDeviceMaps: |tm_*| are global slots another warp of this CTA filled before the call
  // (gemm_dtm_kernels.cu); their users acquire them after the setup barrier below.
  if (!kDeviceMaps && warp == 0 && lane == 0) {
    tma_prefetch_desc(tm_a);
    tma_prefetch_desc(tm_b);
    tma_prefetch_desc(tm_c);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < Cfg::kStages; ++s) {
      // One arrival: the pair leader's expect_tx covering BOTH CTAs' loads (the peer's
      // TMA complete_tx lands on the leader's barrier; no peer arrive is needed since the
      // peer only refills a slot after the multicast commit freed it in both CTAs).
      mbar_init(bar_full + 8 * s, 1);
      mbar_init(bar_empty + 8 * s, 1); // one tcgen05.commit
    }
    for (int a = 0; a < Cfg::kAccStages; ++a) {
      mbar_init(bar_tfull + 8 * a, 1);       // one tcgen05.commit
      mbar_init(bar_tempty + 8 * a, 4 * G);  // one arrive per epilogue warp per CTA
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc<G>(tmem_slot, Cfg::kTmemCols);
  }
  tcgen05_fence_before();
  if constexpr (G == 2) {
    cluster_sync();
  } else {
    __syncthreads();
  }
  tcgen05_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_gen + (tmem_slot - smem_base));

  if (warp == 0) {
    //===------------------------------------------------------------------===//
    // TMA producer
    //===------------------------------------------------------------------===//
    if (!diag_no_load && elect_one()) {
      if constexpr (kDeviceMaps) {
        tensormap_fence_acquire(tm_a);
        tensormap_fence_acquire(tm_b);
      }
      // All loads of the pair complete on the LEADER's full barrier.
      const uint32_t full0 = (G == 2) ? map_to_cta(bar_full, 0) : bar_full;
      const uint64_t policy_keep = l2_policy_evict_last(), policy_stream = l2_policy_evict_first();
      uint32_t stage = 0, phase = 0;
      for (uint32_t tile = cluster_idx; tile < total_tiles; tile += cluster_cnt) {
        const TileCoord tc = decode_tile<Cfg>(tile, tiles_m, tiles_n, group_rows);
        const int m0 = tc.m_blk * Cfg::kTileM + (int)cta_rank * Cfg::kBlockM;
        const int n0 = tc.n_blk * Cfg::kTileN + (int)cta_rank * Cfg::kBlockN;
        for (uint32_t kb = 0; kb < num_kb; ++kb) {
          mbar_wait(bar_empty + 8 * stage, phase ^ 1);
          const uint32_t full = full0 + 8 * stage;
          const int k0 = kb * Cfg::kBlockK;
          if (is_leader) {
            mbar_expect_tx(bar_full + 8 * stage, Cfg::kStageBytes * G);
          }
          if (diag_l2_hint) {
            tma_load_3d_hint<G>(smem_a + stage * Cfg::kABytes, tm_a, full, k0, m0, tc.batch,
                                policy_keep);
            if constexpr (Cfg::kBMajorK) {
              tma_load_3d_hint<G>(smem_b + stage * Cfg::kBBytes, tm_b, full, k0, n0, tc.batch,
                                  policy_stream);
            } else {
#pragma unroll
              for (int j = 0; j < Cfg::kBBoxes; ++j) {
                tma_load_3d_hint<G>(smem_b + stage * Cfg::kBBytes + j * Cfg::kBBoxBytes, tm_b, full,
                                    n0 + j * Cfg::kBBoxN, k0, tc.batch, policy_stream);
              }
            }
          } else {
          if constexpr (Cfg::kTiled) {
            // Coordinates (8 rows x k0, k1, r0 / 8, r1): K tile kb * 8, row tile (row / R0).
            tma_load_4d<G>(smem_a + stage * Cfg::kABytes, tm_a, full, 0, (int)kb * 8, 0, m0 / tile_m0);
            tma_load_4d<G>(smem_b + stage * Cfg::kBBytes, tm_b, full, 0, (int)kb * 8, 0, n0 / tile_n0);
          } else {
          tma_load_3d<G>(smem_a + stage * Cfg::kABytes, tm_a, full, k0, m0, tc.batch);
          if constexpr (Cfg::kBMajorK) {
            tma_load_3d<G>(smem_b + stage * Cfg::kBBytes, tm_b, full, k0, n0, tc.batch);
          } else {
#pragma unroll
            for (int j = 0; j < Cfg::kBBoxes; ++j) {
              tma_load_3d<G>(smem_b + stage * Cfg::kBBytes + j * Cfg::kBBoxBytes, tm_b, full,
                             n0 + j * Cfg::kBBoxN, k0, tc.batch);
            }
          }
          }
          }
          if (++stage == Cfg::kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    //===------------------------------------------------------------------===//
    // MMA issuer (pair leader only)
    //===------------------------------------------------------------------===//
    if (is_leader && elect_one()) {
      uint32_t stage = 0, phase = 0;
      uint32_t acc = 0, acc_phase = 0;
      for (uint32_t tile = cluster_idx; tile < total_tiles; tile += cluster_cnt) {
        mbar_wait(bar_tempty + 8 * acc, acc_phase ^ 1);
        tcgen05_fence_after();
        const uint32_t tmem_d = tmem_base + acc * Cfg::kTileN;
        for (uint32_t kb = 0; kb < num_kb; ++kb) {
          if (!diag_no_load) mbar_wait(bar_full + 8 * stage, phase);
          tcgen05_fence_after();
          const uint32_t a_addr = smem_a + stage * Cfg::kABytes;
          const uint32_t b_addr = smem_b + stage * Cfg::kBBytes;
#pragma unroll
          for (int k = 0; k < Cfg::kKSteps; ++k) {
            // A: K-major SW128 atom rows of 128 B; 8-row groups 1024 B apart.
            uint64_t da, db;
            if constexpr (Cfg::kTiled) {
              // Core matrices 128 B apart along K (one UMMA = 2 of them), 8-row groups 1024 B.
              da = make_smem_desc_interleaved(a_addr + k * 256, 128, 1024);
              db = make_smem_desc_interleaved(b_addr + k * 256, 128, 1024);
            } else {
            da = make_smem_desc_sw128(a_addr + k * 32, 16, 1024);
            if constexpr (Cfg::kBMajorK) {
              db = make_smem_desc_sw128(b_addr + k * 32, 16, 1024);
            } else {
              // MN-major: 128 B (of N) x 8 (K rows) atoms; K groups 1024 B apart,
              // N blocks one TMA box apart; one UMMA consumes kUmmaK rows.
              db = make_smem_desc_sw128(b_addr + k * (Cfg::kUmmaK * 128), Cfg::kBBoxBytes, 1024);
            }
            }
            if (!diag_no_mma) umma_ss<G, Cfg::kKind>(tmem_d, da, db, Cfg::kIdesc, (kb | k) != 0);
          }
          umma_commit<G>(bar_empty + 8 * stage);  // smem slot free once these MMAs retire
          if (++stage == Cfg::kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit<G>(bar_tfull + 8 * acc);  // accumulator ready for both CTAs' epilogues
        if (++acc == Cfg::kAccStages) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
  } else if (warp >= 4) {
    //===------------------------------------------------------------------===//
    // Epilogue: TMEM -> registers -> swizzled smem -> TMA store
    //===------------------------------------------------------------------===//
    const uint32_t ew = warp - 4;          // == warp % 4: TMEM lane quarter
    const uint32_t row = ew * 32 + lane;   // row within this CTA's 128
    const bool epi_leader = (warp == 4) && (lane == 0);
    const uint32_t tempty0 = (G == 2) ? map_to_cta(bar_tempty, 0) : bar_tempty;
    const bool accumulate = (flags & 1) != 0;
    const bool relu = (flags & 2) != 0;  // kBias exports only
    (void)relu;
    // The fused consumer: x + bias[column], then optionally max(x, 0) with arith.maximumf /
    // arith.maxsi semantics (NaN propagates, -0 -> +0).
    auto fuse = [&](uint32_t bits, int gn) -> uint32_t {
      if constexpr (!Cfg::kBias) {
        return bits;
      } else if constexpr (Cfg::kKind == KIND_I8) {
        int32_t x = (int32_t)bits;
        if (gn < N) x += __ldg(reinterpret_cast<const int32_t*>(bias) + gn);
        if (relu) x = max(x, 0);
        return (uint32_t)x;
      } else {
        float x = __uint_as_float(bits);
        if (gn < N) x += __ldg(reinterpret_cast<const float*>(bias) + gn);
        if (relu) x = (x > 0.f || x != x) ? x : 0.f;
        return __float_as_uint(x);
      }
    };
    uint32_t acc = 0, acc_phase = 0;
    uint32_t buf = 0;
    if constexpr (kDeviceMaps) {
      if (epi_leader) tensormap_fence_acquire(tm_c);
    }
    for (uint32_t tile = cluster_idx; tile < total_tiles; tile += cluster_cnt) {
      const TileCoord tc = decode_tile<Cfg>(tile, tiles_m, tiles_n, group_rows);
      const int m0 = tc.m_blk * Cfg::kTileM + (int)cta_rank * Cfg::kBlockM;
      const int n0 = tc.n_blk * Cfg::kTileN;
      mbar_wait(bar_tfull + 8 * acc, acc_phase);
      tcgen05_fence_after();
      const uint32_t taddr0 = tmem_base + acc * Cfg::kTileN + ((ew * 32u) << 16);
#pragma unroll 1
      for (int c = 0; c < Cfg::kTileN / Cfg::kEpiCols; ++c) {
        // Buffer `buf` was last read by the TMA store issued two chunks ago.
        if (epi_leader) tma_store_wait_read<Cfg::kEpiBufs - 1>();
        named_barrier_sync(1, 128);
        uint32_t v[32];
        if constexpr (Cfg::kOut16) {
          // 64 accumulator columns -> 32 packed 16-bit pairs (round to nearest even).
          uint32_t hi[32];
          tmem_ld_32x32b_x32(taddr0 + c * Cfg::kEpiCols, v);
          tmem_ld_32x32b_x32(taddr0 + c * Cfg::kEpiCols + 32, hi);
          tmem_wait_ld();
          if constexpr (Cfg::kBias) {
            const int gn = n0 + c * Cfg::kEpiCols;
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              v[i] = fuse(v[i], gn + i);
              hi[i] = fuse(hi[i], gn + 32 + i);
            }
          }
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            v[i] = pack_pair<Cfg::kAbFmt == FMT_BF16>(__uint_as_float(v[2 * i]),
                                                       __uint_as_float(v[2 * i + 1]));
          }
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            v[16 + i] = pack_pair<Cfg::kAbFmt == FMT_BF16>(__uint_as_float(hi[2 * i]),
                                                            __uint_as_float(hi[2 * i + 1]));
          }
        } else {
          tmem_ld_32x32b_x32(taddr0 + c * Cfg::kEpiCols, v);
          tmem_wait_ld();
        }
        if (c == Cfg::kTileN / Cfg::kEpiCols - 1) {
          // Whole accumulator stage has been read out: hand it back to the MMA warp.
          tcgen05_fence_before();
          __syncwarp();
          if (lane == 0) {
            if constexpr (G == 2) {
              mbar_arrive_cluster(tempty0 + 8 * acc);
            } else {
              mbar_arrive(bar_tempty + 8 * acc);
            }
          }
        }
        if (accumulate && !Cfg::kOut16) {
          const int gm = m0 + (int)row;
          const int gn = n0 + c * Cfg::kEpiCols;
          if (Cfg::kTiled && gm < M) {
            // Result tiles [M1, N1, M0, N0]: element (gm, gn + i).
            const size_t row_base =
                (size_t)(gm / tile_m0) * ((size_t)N * tile_m0) + (size_t)(gm % tile_m0) * tile_n0;
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              const int n = gn + i;
              if (n >= N) continue;
              const size_t at = row_base + (size_t)(n / tile_n0) * ((size_t)tile_m0 * tile_n0) + n % tile_n0;
              if constexpr (Cfg::kKind == KIND_I8) {
                v[i] = (uint32_t)((int32_t)v[i] + reinterpret_cast<const int32_t*>(out)[at]);
              } else {
                v[i] = __float_as_uint(__uint_as_float(v[i]) + reinterpret_cast<const float*>(out)[at]);
              }
            }
          } else if (gm < M) {
            const size_t base = ((size_t)tc.batch * M + gm) * (size_t)N + gn;
            if constexpr (Cfg::kKind == KIND_I8) {
              const int32_t* cin = reinterpret_cast<const int32_t*>(out) + base;
#pragma unroll
              for (int i = 0; i < 32; ++i)
                if (gn + i < N) v[i] = (uint32_t)((int32_t)v[i] + cin[i]);
            } else {
              const float* cin = reinterpret_cast<const float*>(out) + base;
#pragma unroll
              for (int i = 0; i < 32; ++i)
                if (gn + i < N) v[i] = __float_as_uint(__uint_as_float(v[i]) + cin[i]);
            }
          }
        }
        if constexpr (Cfg::kBias && !Cfg::kOut16) {
          const int gn = n0 + c * Cfg::kEpiCols;
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] = fuse(v[i], gn + i);
        }
        // 128-byte rows, 16-byte chunks XOR-swizzled by (row % 8) == SWIZZLE_128B.
        const uint32_t row_addr = smem_epi + buf * Cfg::kEpiBytes + row * 128;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const uint32_t dst = row_addr + ((i ^ (row & (Cfg::kTiled ? epi_swizzle : 7u))) << 4);
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "r"(v[4 * i]),
                       "r"(v[4 * i + 1]), "r"(v[4 * i + 2]), "r"(v[4 * i + 3])
                       : "memory");
        }
        fence_proxy_async_smem();
        named_barrier_sync(1, 128);
        if (epi_leader && !diag_no_store) {
          if constexpr (Cfg::kTiled) {
            // Coordinates (n0, n1, m0, m1) of the result tiles.
            const int col = n0 + c * Cfg::kEpiCols;
            tma_store_4d(tm_c, smem_epi + buf * Cfg::kEpiBytes, col % tile_n0, col / tile_n0, 0,
                         m0 / tile_m0);
          } else {
            tma_store_3d(tm_c, smem_epi + buf * Cfg::kEpiBytes, n0 + c * Cfg::kEpiCols, m0,
                         tc.batch);
          }
          tma_store_commit();
        }
        buf ^= 1;
      }
      if (++acc == Cfg::kAccStages) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
    if (epi_leader) tma_store_wait<0>();
  }

  // Warps 0 and 1 ran their loops on one elected lane; reconverge before the .aligned
  // barrier / dealloc instructions below.
  __syncwarp();

  // Teardown: nobody may free TMEM (or exit while the peer's MMA still reads our smem)
  // until every role of every CTA in the cluster is done.
  tcgen05_fence_before();
  if constexpr (G == 2) {
    cluster_sync();
  } else {
    __syncthreads();
  }
  if (warp == 2) {
    tcgen05_fence_after();
    tmem_dealloc<G>(tmem_base, Cfg::kTmemCols);
  }
}

// Fills the `<kernel>__tma` recipe for a GEMM config. Push constants: 0=M 1=N 2=K
// 3=batch 4=flags. Bindings: 0=A 1=B 2=C (bias exports: 2=bias 3=C).
template <class Cfg>
constexpr iree_b200_tma_recipe_t make_gemm_recipe() {
  iree_b200_tma_recipe_t r = {};
  r.magic = IREE_B200_TMA_RECIPE_MAGIC;
  r.map_count = 3;
  const uint32_t ab_type = Cfg::kKind != KIND_F16
                               ? IREE_B200_TMA_U8
                               : (Cfg::kAbFmt == FMT_BF16 ? IREE_B200_TMA_BF16 : IREE_B200_TMA_F16);
  // A [batch, M, K]
  r.maps[0].binding = 0;
  r.maps[0].elem_type = ab_type;
  r.maps[0].elem_size = Cfg::kElemSize;
  r.maps[0].rank = 3;
  r.maps[0].dim[0] = IREE_B200_TMA_DIM_CONST(2);
  r.maps[0].dim[1] = IREE_B200_TMA_DIM_CONST(0);
  r.maps[0].dim[2] = IREE_B200_TMA_DIM_CONST(3);
  r.maps[0].box[0] = Cfg::kBlockK;
  r.maps[0].box[1] = Cfg::kBlockM;
  r.maps[0].box[2] = 1;
  r.maps[0].swizzle = IREE_B200_TMA_SWIZZLE_128B;
  // B [batch, K, N] or [batch, N, K]
  r.maps[1].binding = 1;
  r.maps[1].elem_type = ab_type;
  r.maps[1].elem_size = Cfg::kElemSize;
  r.maps[1].rank = 3;
  if (Cfg::kBMajorK) {
    r.maps[1].dim[0] = IREE_B200_TMA_DIM_CONST(2);
    r.maps[1].dim[1] = IREE_B200_TMA_DIM_CONST(1);
    r.maps[1].box[0] = Cfg::kBlockK;
    r.maps[1].box[1] = Cfg::kBlockN;
  } else {
    r.maps[1].dim[0] = IREE_B200_TMA_DIM_CONST(1);
    r.maps[1].dim[1] = IREE_B200_TMA_DIM_CONST(2);
    r.maps[1].box[0] = Cfg::kBBoxN;
    r.maps[1].box[1] = Cfg::kBlockK;
  }
  r.maps[1].dim[2] = IREE_B200_TMA_DIM_CONST(3);
  r.maps[1].box[2] = 1;
  r.maps[1].swizzle = IREE_B200_TMA_SWIZZLE_128B;
  // C [batch, M, N] (32-bit elements)
  r.maps[2].binding = Cfg::kBias ? 3 : 2;
  r.maps[2].elem_type = Cfg::kOut16 ? ab_type
                                    : (Cfg::kKind == KIND_I8 ? IREE_B200_TMA_I32 : IREE_B200_TMA_F32);
  r.maps[2].elem_size = Cfg::kOut16 ? 2 : 4;
  r.maps[2].rank = 3;
  r.maps[2].dim[0] = IREE_B200_TMA_DIM_CONST(1);
  r.maps[2].dim[1] = IREE_B200_TMA_DIM_CONST(0);
  r.maps[2].dim[2] = IREE_B200_TMA_DIM_CONST(3);
  r.maps[2].box[0] = Cfg::kEpiCols;
  r.maps[2].box[1] = Cfg::kBlockM;
  r.maps[2].box[2] = 1;
  r.maps[2].swizzle = IREE_B200_TMA_SWIZZLE_128B;
  return r;
}

// `<kernel>__tma` of the tiled (linalg.mmt4d) exports. Bindings 0 lhs [M1,K1,M0,K0], 1 rhs
// [N1,K1,N0,K0], 2 out [M1,N1,M0,N0]. Push constants: 0 M1, 1 N1, 2 K1, 3 M0, 4 N0, 5 K0,
// 6 flags, 7 M0 / 8, 8 = 128 / M0, 9 N0 / 8, 10 = kBlockN / N0, 11 = min(N0, 32),
// 12 = 32 / min(N0, 32), 13 = swizzle mode of the result map (min(N0, 32) * 4 bytes wide).
//
// Operands: K0 rows are 16 bytes and tile rows are multiples of 8, so 8 tile rows x K0 are
// 128 contiguous bytes = one UMMA core matrix. The map sees the tiled tensor [R1][K1][R0/8]
// [8 x K0] as (8 x K0, K1, R0 / 8, R1) — its dense order is (8 x K0, R0 / 8, K1, R1) — and a
// box of (8 x K0, 8, R0 / 8, rows / R0) lands UNSWIZZLED as [8-row group][8 K tiles][128 B]:
// the interleaved K-major UMMA layout with core matrices 128 B apart along K and 1024 B
// apart along the rows. (A 16-byte inner box under SWIZZLE_128B, the other way to reach the
// row-major GEMM's tile, faults on the hardware: tools/tma4d_probe.cu.)
// Result: (n0, n1, m0, m1) with boxes of 32 columns; the swizzle span equals the box row.
template <class Cfg>
constexpr iree_b200_tma_recipe_t make_gemm_tiled_recipe() {
  iree_b200_tma_recipe_t r = {};
  r.magic = IREE_B200_TMA_RECIPE_MAGIC;
  r.map_count = 3;
  const uint32_t ab_type = Cfg::kKind != KIND_F16
                               ? IREE_B200_TMA_U8
                               : (Cfg::kAbFmt == FMT_BF16 ? IREE_B200_TMA_BF16 : IREE_B200_TMA_F16);
  for (int i = 0; i < 2; ++i) {
    r.maps[i].binding = i;
    r.maps[i].elem_type = ab_type;
    r.maps[i].elem_size = Cfg::kElemSize;
    r.maps[i].rank = 4;
    r.maps[i].dim[0] = 128 / Cfg::kElemSize;                     // 8 rows x k0
    r.maps[i].dim[1] = IREE_B200_TMA_DIM_CONST(2);               // k1
    r.maps[i].dim[2] = IREE_B200_TMA_DIM_CONST(i == 0 ? 7 : 9);  // r0 / 8
    r.maps[i].dim[3] = IREE_B200_TMA_DIM_CONST(i == 0 ? 0 : 1);  // r1
    r.maps[i].box[0] = 128 / Cfg::kElemSize;
    r.maps[i].box[1] = 8;
    r.maps[i].box[2] = IREE_B200_TMA_DIM_CONST(i == 0 ? 7 : 9);
    r.maps[i].box[3] = IREE_B200_TMA_DIM_CONST(i == 0 ? 8 : 10);
    r.maps[i].swizzle = IREE_B200_TMA_SWIZZLE_NONE;
    r.maps[i].order = IREE_B200_TMA_ORDER(0, 2, 1, 3);
  }
  r.maps[2].binding = 2;
  r.maps[2].elem_type = Cfg::kKind == KIND_I8 ? IREE_B200_TMA_I32 : IREE_B200_TMA_F32;
  r.maps[2].elem_size = 4;
  r.maps[2].rank = 4;
  r.maps[2].dim[0] = IREE_B200_TMA_DIM_CONST(4);  // n0
  r.maps[2].dim[1] = IREE_B200_TMA_DIM_CONST(1);  // n1
  r.maps[2].dim[2] = IREE_B200_TMA_DIM_CONST(3);  // m0
  r.maps[2].dim[3] = IREE_B200_TMA_DIM_CONST(0);  // m1
  r.maps[2].box[0] = IREE_B200_TMA_DIM_CONST(11);
  r.maps[2].box[1] = IREE_B200_TMA_DIM_CONST(12);
  r.maps[2].box[2] = IREE_B200_TMA_DIM_CONST(3);
  r.maps[2].box[3] = IREE_B200_TMA_DIM_CONST(8);
  r.maps[2].swizzle = IREE_B200_TMA_DIM_CONST(13);
  r.maps[2].order = IREE_B200_TMA_ORDER(0, 2, 1, 3);
  return r;
}

}  // namespace b200
