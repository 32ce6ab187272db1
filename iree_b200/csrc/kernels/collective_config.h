// collective_config.h — geometry of the symmetric-region collective kernels
// (host-includable; shared by collective_kernels.cu, the manifest tool and hal/channel.cc).

#pragma once

#include <stdint.h>

namespace b200 {

struct CollectiveConfig {
  static constexpr int kMaxWorld = 8;    // one NVSwitch domain of B200s
  static constexpr int kMaxBlocks = 64;  // flag words: [kMaxBlocks][kMaxWorld] u32 per rank
  static constexpr int kThreads = 1024;
  // Defaults of IREE_B200_ALL_REDUCE_TUNING (profiles/tp_sweep_n8_r2.jsonl: with <= 24 CTAs the
  // software-pipelined form with 2 requests per thread is the fastest or level with it).
  static constexpr int kUnroll = 2;      // 16-byte multimem requests in flight per thread
  static constexpr int kPipelined = 1;   // loads of step i+1 issued before the stores of step i
  static constexpr int kFlagBytes = kMaxBlocks * kMaxWorld * 4;
  static constexpr int kDefaultBlocks = 16;
  // Device table: bases[kMaxWorld], flag_blocks[kMaxWorld] (u64 each), then
  // {u32 ready, u32 status, u64 watchdog_cycles}.
  static constexpr int kTableBytes = 2 * kMaxWorld * 8 + 16;
  static constexpr int kReadyOffset = 2 * kMaxWorld * 8;
  static constexpr int kStatusOffset = kReadyOffset + 4;
  static constexpr int kWatchdogOffset = kReadyOffset + 8;
  // Status word values (low byte) left by a watchdog expiry.
  static constexpr uint32_t kStatusPutStuck = 1, kStatusWaitStuck = 2, kStatusNeverReady = 3;
};

}  // namespace b200
