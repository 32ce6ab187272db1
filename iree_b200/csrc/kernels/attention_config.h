// attention_config.h — compile-time geometry of the attention kernels (host-includable).
//
// Shared by the kernels (attention_sm100.cuh, attention_kernels.cu) and the manifest tool
// that writes the export table of the CDA1 executable.

#pragma once

#include <stdint.h>

namespace b200 {

// Prefill kernel: one CTA owns two 128-row Q tiles of one (batch*head) slice and walks
// K2 in blocks of 128 rows. Head dimension (K1 == N) is fixed per export.
template <int kHeadDim_>
struct AttnPrefillConfig {
  static constexpr int kHeadDim = kHeadDim_;       // K1 == N
  static constexpr int kBlockM = 128;              // rows per Q tile (= TMEM lanes)
  static constexpr int kTilesPerCta = 2;           // ping-pong Q tiles
  static constexpr int kBlockN = 128;              // K2 rows per block
  static constexpr int kKStages = 2;
  static constexpr int kVStages = 2;
  static constexpr int kChunkElems = 64;           // 128-byte swizzle atom row (16-bit elems)
  static constexpr int kChunks = kHeadDim / kChunkElems;
  static constexpr int kChunkBytes = 128 * 128;    // [128 rows x 128 B]
  static constexpr int kTileBytes = kChunks * kChunkBytes;  // Q tile == K block == V block
  static constexpr int kSmemTiles = (kTilesPerCta + kKStages + kVStages) * kTileBytes;
  static constexpr int kNumBarriers = 2 + 2 + 2 * kKStages + 2 * kVStages + 2 + 4 + 1;
  static constexpr int kSmemBytes = 1024 + kSmemTiles + kNumBarriers * 8 + 16;
  static constexpr int kThreads = 384;  // warps 0-3: TMA / MMA / TMEM alloc / spare; 4-11 softmax
  static constexpr int kTmemCols = 512;  // S0 | S1 | O0 | O1, 128 f32 columns each
  static_assert(kHeadDim == 128, "only head_dim 128 is instantiated");
  static_assert(kSmemBytes <= 232448, "exceeds 227 KiB of shared memory");
};

// Decode (split-KV) kernels: K1 == N == 128, one CTA per (split, m, batch).
struct AttnDecodeConfig {
  static constexpr int kHeadDim = 128;
  static constexpr int kThreads = 256;
  static constexpr int kChunkRows = 256;  // K2 rows per CTA
  // Workspace record per (batch, m, split): kHeadDim unnormalised f32 sums, max, sum.
  static constexpr int kRecordFloats = kHeadDim + 2;
};

// Decode, single-dispatch form: a cluster of kSplits CTAs owns one (batch, m) row; each CTA
// streams its share of K and V through a bulk-copy (TMA 1-D) ring in shared memory and
// the cluster merges the partial softmax states through distributed shared memory.
struct AttnDecodeClusterConfig {
  static constexpr int kHeadDim = 128;
  static constexpr int kSplits = 8;          // == compiled-in cluster size (portable maximum)
  static constexpr int kStageRows = 32;      // K2 rows per ring stage
  static constexpr int kStages = 6;
  static constexpr int kConsumerWarps = 8;   // 16 half-warps, 2 rows of a stage each
  static constexpr int kThreads = (kConsumerWarps + 1) * 32;  // + 1 producer warp
  static constexpr int kRowBytes = kHeadDim * 2;
  static constexpr int kStageBytes = 2 * kStageRows * kRowBytes;  // K rows then V rows
  static constexpr int kRecordFloats = kHeadDim + 2;              // sums, max, sum
  static constexpr int kGroups = kConsumerWarps * 2;
  static constexpr int kSmemBytes = 128 /*alignment slack*/ + kStages * kStageBytes +
                                    (kGroups + kSplits) * kRecordFloats * 4 + 2 * kStages * 8 + 64;
  static constexpr int kMinRowsPerSplit = 32;  // selection guard: K2 >= kSplits * this
  static_assert(kStageRows == kGroups * 2, "two rows per half-warp and stage");
};

// Generic SIMT kernel limits (matcher guards).
struct AttnGenericConfig {
  static constexpr int kThreads = 128;
  static constexpr int kMaxK1 = 512;
  static constexpr int kMaxN = 512;
};

}  // namespace b200
