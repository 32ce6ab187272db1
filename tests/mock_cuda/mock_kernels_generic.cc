// Host execution of generic_kernels.cu for the mock driver: mock_run_<kernel>(grid, block,
// params) unpacks the ABI parameters exactly as the GPU would see them (one slot per binding
// pointer, then one per 32-bit constant) and runs the kernel body on OS threads.
#include "shim_common.h"

#include "../../iree_b200/csrc/kernels/generic_kernels.cu"

#define MOCK_RUN(NAME, T, ACC)                                                                  \
  extern "C" __attribute__((visibility("default"))) int mock_run_##NAME(                         \
      const unsigned grid[3], const unsigned block[3], void** p) {                               \
    if (block[0] != 256) return 1;                                                               \
    run_grid(grid, [&] {                                                                         \
      NAME(arg<const T*>(p, 0), arg<const T*>(p, 1), arg<ACC*>(p, 2), arg<int>(p, 3),            \
           arg<int>(p, 4), arg<int>(p, 5), arg<int>(p, 6), arg<int>(p, 7));                      \
    });                                                                                          \
    return 0;                                                                                    \
  }
#define MOCK_RUN_BIAS(NAME, T, ACC)                                                             \
  extern "C" __attribute__((visibility("default"))) int mock_run_##NAME(                         \
      const unsigned grid[3], const unsigned block[3], void** p) {                               \
    if (block[0] != 256) return 1;                                                               \
    run_grid(grid, [&] {                                                                         \
      NAME(arg<const T*>(p, 0), arg<const T*>(p, 1), arg<const ACC*>(p, 2), arg<ACC*>(p, 3),     \
           arg<int>(p, 4), arg<int>(p, 5), arg<int>(p, 6), arg<int>(p, 7), arg<int>(p, 8));      \
    });                                                                                          \
    return 0;                                                                                    \
  }
MOCK_RUN(iree_b200_matmul_generic_bf16_f32, __nv_bfloat16, float)
MOCK_RUN(iree_b200_matmul_generic_f16_f32, __half, float)
MOCK_RUN(iree_b200_matmul_generic_i8_i32, int8_t, int32_t)
MOCK_RUN(iree_b200_matmul_generic_f8e4m3fn_f32, F8E4M3FN, float)
MOCK_RUN(iree_b200_matmul_generic_f8e5m2_f32, F8E5M2, float)
MOCK_RUN_BIAS(iree_b200_matmul_generic_bias_bf16_f32, __nv_bfloat16, float)
MOCK_RUN_BIAS(iree_b200_matmul_generic_bias_f16_f32, __half, float)
MOCK_RUN_BIAS(iree_b200_matmul_generic_bias_i8_i32, int8_t, int32_t)
