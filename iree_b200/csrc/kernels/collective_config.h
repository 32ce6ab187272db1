// collective_config.h — geometry of the symmetric-region collective kernels
// (host-includable; shared by collective_kernels.cu, the manifest tool and hal/channel.cc).

#pragma once

#include <stdint.h>

namespace b200 {

struct CollectiveConfig {
  static constexpr int kMaxWorld = 8;    // one NVSwitch domain of B200s
  static constexpr int kMaxBlocks = 64;  // flag words: [kMaxBlocks][kMaxWorld] u32 per rank
  static constexpr int kThreads = 1024;
  static constexpr int kUnroll = 4;      // 16-byte multimem requests in flight per thread
  static constexpr int kFlagBytes = kMaxBlocks * kMaxWorld * 4;
  static constexpr int kDefaultBlocks = 16;
};

}  // namespace b200
