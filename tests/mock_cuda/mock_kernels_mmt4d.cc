// Host execution of mmt4d_kernels.cu for the mock driver (see mock_kernels_generic.cc).
#include "shim_common.h"

#include "../../iree_b200/csrc/kernels/mmt4d_kernels.cu"

#define MOCK_RUN_MMT4D(NAME, T, ACC)                                                            \
  extern "C" __attribute__((visibility("default"))) int mock_run_##NAME(                         \
      const unsigned grid[3], const unsigned block[3], void** p) {                               \
    if (block[0] != 256) return 1;                                                               \
    run_grid(grid, [&] {                                                                         \
      NAME(arg<const T*>(p, 0), arg<const T*>(p, 1), arg<ACC*>(p, 2), arg<int>(p, 3),            \
           arg<int>(p, 4), arg<int>(p, 5), arg<int>(p, 6), arg<int>(p, 7), arg<int>(p, 8),       \
           arg<int>(p, 9));                                                                      \
    });                                                                                          \
    return 0;                                                                                    \
  }
MOCK_RUN_MMT4D(iree_b200_mmt4d_generic_bf16_f32, __nv_bfloat16, float)
MOCK_RUN_MMT4D(iree_b200_mmt4d_generic_f16_f32, __half, float)
MOCK_RUN_MMT4D(iree_b200_mmt4d_generic_i8_i32, int8_t, int32_t)
MOCK_RUN_MMT4D(iree_b200_mmt4d_generic_f32_f32, float, float)

#define MOCK_RUN_RELAYOUT(NAME, T)                                                              \
  extern "C" __attribute__((visibility("default"))) int mock_run_##NAME(                         \
      const unsigned grid[3], const unsigned block[3], void** p) {                               \
    if (block[0] != 256) return 1;                                                               \
    run_grid(grid, [&] {                                                                         \
      NAME(arg<const T*>(p, 0), arg<T*>(p, 1), arg<int>(p, 2), arg<int>(p, 3), arg<int>(p, 4),   \
           arg<int>(p, 5), arg<int>(p, 6));                                                      \
    });                                                                                          \
    return 0;                                                                                    \
  }
MOCK_RUN_RELAYOUT(iree_b200_mmt4d_relayout_b8, uint8_t)
MOCK_RUN_RELAYOUT(iree_b200_mmt4d_relayout_b16, uint16_t)
MOCK_RUN_RELAYOUT(iree_b200_mmt4d_relayout_b32, uint32_t)
