// attention_config.h — compile-time geometry of the attention kernels (host-includable).

#pragma once

#include <stdint.h>

namespace b200 {}  // namespace b200
