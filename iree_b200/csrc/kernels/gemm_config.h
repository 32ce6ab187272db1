// gemm_config.h — compile-time geometry of the tcgen05 GEMM family.
//
// Pure constexpr arithmetic with no CUDA dependency, so the same definitions feed the
// kernels (gemm_sm100.cuh) and the host-side manifest tool that writes the export table
// of the CDA1 executable (block dims, dynamic shared memory, constant/binding counts —
// what compiler/plugins/target/CUDA/CUDATarget.cpp:661-691 takes from the export op).

#pragma once

#include <stdint.h>

#if defined(__CUDACC__)
#define B200_HOST_DEVICE __host__ __device__
#else
#define B200_HOST_DEVICE
#endif

namespace b200 {

// MMA kinds.
enum : int { KIND_F16 = 0, KIND_I8 = 1, KIND_F8 = 2 /* tcgen05 kind::f8f6f4, 8-bit operands */ };

enum : uint32_t { FMT_F16 = 0, FMT_BF16 = 1, FMT_U8 = 0, FMT_S8 = 1, FMT_E4M3 = 0, FMT_E5M2 = 1 };
enum : uint32_t { CFMT_F16 = 0, CFMT_F32 = 1, CFMT_S32 = 2 };
enum : uint32_t { MAJOR_K = 0, MAJOR_MN = 1 };

// 32-bit instruction descriptor:
//   [4,6) D format  [7,10) A format  [10,13) B format  [15] A major  [16] B major
//   [17,23) N>>3    [24,29) M>>4
B200_HOST_DEVICE constexpr uint32_t make_idesc(uint32_t c_fmt, uint32_t a_fmt, uint32_t b_fmt,
                                                  uint32_t a_major, uint32_t b_major, uint32_t m,
                                                  uint32_t n) {
  return (c_fmt << 4) | (a_fmt << 7) | (b_fmt << 10) | (a_major << 15) | (b_major << 16) |
         ((n >> 3) << 17) | ((m >> 4) << 24);
}

// kOut16: the epilogue truncates the f32 accumulator to the operand type (the fused
// `linalg.matmul + arith.truncf` dispatch), halving the C traffic.
// kBias: the dispatch also holds the consumer elementwise op `act(x + bias[n])` (a bias
// binding between rhs and out; see gemm_bias_kernels.cu).
// kTiled: linalg.mmt4d — operands and result are data-tiled ([R1, K1, R0, K0] with K0 * size =
// 16 bytes; result [M1, N1, M0, N0]) and reach the same shared-memory tiles through rank-4
// tensor maps that walk the tiled tensor in (k0, k1, r0, r1) order (gemm_sm100.cuh).
template <int kCtaGroup_, int kKind_, uint32_t kAbFmt_, bool kBMajorK_, int kStages_,
          bool kOut16_ = false, bool kBias_ = false, bool kTiled_ = false>
struct GemmConfig {
  static constexpr bool kOut16 = kOut16_;
  static constexpr bool kBias = kBias_;
  static constexpr bool kTiled = kTiled_;
  static constexpr int kCtaGroup = kCtaGroup_;
  static constexpr int kKind = kKind_;
  static constexpr uint32_t kAbFmt = kAbFmt_;
  static constexpr bool kBMajorK = kBMajorK_;
  static constexpr int kStages = kStages_;

  static constexpr int kElemSize = (kKind == KIND_F16) ? 2 : 1;
  static constexpr int kBlockM = 128;                    // rows per CTA
  static constexpr int kTileM = kBlockM * kCtaGroup;     // rows per cluster tile
  static constexpr int kTileN = 256;                     // UMMA N
  static constexpr int kBlockN = kTileN / kCtaGroup;     // B columns loaded per CTA
  static constexpr int kBlockK = 128 / kElemSize;        // elements per 128-byte row
  static constexpr int kUmmaK = 32 / kElemSize;
  static constexpr int kKSteps = kBlockK / kUmmaK;       // 4

  static constexpr int kABytes = kBlockM * 128;                 // 16 KiB
  static constexpr int kBBytes = kBlockN * kBlockK * kElemSize; // 16 / 32 KiB
  static constexpr int kStageBytes = kABytes + kBBytes;

  // MN-major B: [kBlockK rows][128 B of N] blocks, one TMA box each.
  static constexpr int kBBoxN = 128 / kElemSize;               // N elements per box
  static constexpr int kBBoxBytes = kBlockK * 128;             // bytes per box
  static constexpr int kBBoxes = kBBytes / kBBoxBytes;

  static constexpr int kAccStages = 2;
  static constexpr int kTmemCols = kAccStages * kTileN;  // 512
  static constexpr int kEpiCols = kOut16 ? 64 : 32;      // columns per epilogue chunk (128 B rows)
  static constexpr int kEpiBytes = kBlockM * 128;        // 16 KiB
  static constexpr int kEpiBufs = 2;

  static constexpr int kSmemTiles = kStages * kStageBytes + kEpiBufs * kEpiBytes;
  static constexpr int kNumBarriers = 2 * kStages + 2 * kAccStages;
  static constexpr int kSmemBytes = 1024 /*align slack*/ + kSmemTiles + kNumBarriers * 8 + 16;

  static constexpr int kThreads = 256;
  static constexpr int kGroupM = 8;  // raster group height in tiles

  static constexpr uint32_t kIdesc =
      make_idesc(kKind == KIND_I8 ? CFMT_S32 : CFMT_F32, kAbFmt, kAbFmt, MAJOR_K,
                 kBMajorK ? MAJOR_K : MAJOR_MN, kTileM, kTileN);

  static_assert(kSmemBytes <= 232448, "exceeds 227 KiB of shared memory");
};

// MXFP4 block-scaled GEMM (mx_gemm_sm100.cuh): tcgen05.mma.kind::mxf4.block_scale.
// kCtaGroup 2 pairs two CTAs of a cluster on a 256 x 256 tile (each CTA: 128 rows of A and
// 128 of the 256 columns of B in its shared memory, all 256 columns' scales in its TMEM).
template <int kCtaGroup_>
struct MxGemmConfig {
  static constexpr int kCtaGroup = kCtaGroup_;
  static constexpr int kBlockM = 128;                     // A rows per CTA
  static constexpr int kTileM = kBlockM * kCtaGroup, kTileN = 256;
  static constexpr int kBlockN = kTileN / kCtaGroup;      // B rows (result columns) loaded per CTA
  static constexpr int kBlockKBytes = 128;                // packed bytes of K per row and stage
  static constexpr int kBlockK = 2 * kBlockKBytes;        // 256 elements
  static constexpr int kUmmaK = 64;                       // elements per MMA
  static constexpr int kKSteps = kBlockK / kUmmaK;        // 4
  static constexpr int kStages = kCtaGroup == 2 ? 5 : 3;
  static constexpr int kABytes = kBlockM * kBlockKBytes;  // 16 KiB
  static constexpr int kBBytes = kBlockN * kBlockKBytes;  // 16 / 32 KiB
  static constexpr int kAtomBytes = 512;                  // 128 rows x 4 scales
  static constexpr int kSfaBytes = 2 * kAtomBytes;        // 2 K atoms of this CTA's 128 rows
  static constexpr int kSfbBytes = 4 * kAtomBytes;        // 2 K atoms x 2 halves of the 256 columns
  static constexpr int kStageBytes = kABytes + kBBytes + kSfaBytes + kSfbBytes;
  // Epilogue: two warpgroups (column halves), one 16 KiB staging buffer each.
  static constexpr int kEpiCols = 32, kEpiBytes = kBlockM * 128, kEpiGroups = 2;
  static constexpr int kSmemTiles = kStages * kStageBytes + kEpiGroups * kEpiBytes;
  static constexpr int kNumBarriers = 3 * kStages + 2;    // full, empty, scales-landed; tfull, tempty
  static constexpr int kSmemBytes = 1024 + kSmemTiles + kNumBarriers * 8 + 16;
  static constexpr int kThreads = 384;                    // warps 0-3 control, 4-11 epilogue
  static constexpr int kTmemCols = 512;
  static constexpr int kSfCol0 = kTileN;                  // first scale column
  static constexpr int kSfColsPerStage = 8 + 16;
  static constexpr int kGroupM = 8;
  static constexpr int kScaleLookahead = 2;               // stages the scale loader runs ahead
  // Block-scaled instruction descriptor (cute/arch/mma_sm100_desc.hpp
  // InstrDescriptorBlockScaled): [7,10) A format, [10,13) B format (kind::mxf4: E2M1 = 1;
  // the value 5 belongs to kind::mxf8f6f4's unpacked form and is illegal here), [15]/[16]
  // majors (K), [17,23) N >> 3, [23] scale format (1 = UE8M0), [24,29) M >> 4; the scale
  // ids ([4,6) B, [29,31) A) are OR-ed in per MMA.
  static constexpr uint32_t kIdesc = (1u << 7) | (1u << 10) | ((kTileN >> 3) << 17) | (1u << 23) |
                                     ((kTileM >> 4) << 24);
  static_assert(kSmemBytes <= 232448, "exceeds 227 KiB of shared memory");
  static_assert(kSfCol0 + kStages * kSfColsPerStage <= kTmemCols, "TMEM columns");
  static_assert(kScaleLookahead < kStages, "lookahead");
};

}  // namespace b200
