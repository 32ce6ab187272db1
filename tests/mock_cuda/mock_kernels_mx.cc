// Host execution of mx_kernels.cu for the mock driver (see mock_kernels_generic.cc).
#include "shim_common.h"

#include "../../iree_b200/csrc/kernels/mx_kernels.cu"

extern "C" __attribute__((visibility("default"))) int mock_run_iree_b200_matmul_generic_mxfp4_f32(
    const unsigned grid[3], const unsigned block[3], void** p) {
  if (block[0] != 256) return 1;
  run_grid(grid, [&] {
    iree_b200_matmul_generic_mxfp4_f32(arg<const uint8_t*>(p, 0), arg<const uint8_t*>(p, 1),
                                       arg<const uint8_t*>(p, 2), arg<const uint8_t*>(p, 3),
                                       arg<float*>(p, 4), arg<int>(p, 5), arg<int>(p, 6),
                                       arg<int>(p, 7), arg<int>(p, 8), arg<int>(p, 9));
  });
  return 0;
}

extern "C" __attribute__((visibility("default"))) int mock_run_iree_b200_matmul_mxfp4_dequant_bf16(
    const unsigned grid[3], const unsigned block[3], void** p) {
  if (block[0] != 256) return 1;
  run_grid(grid, [&] {
    iree_b200_matmul_mxfp4_dequant_bf16(arg<const uint8_t*>(p, 0), arg<const uint8_t*>(p, 1),
                                        arg<uint16_t*>(p, 2), arg<int>(p, 3), arg<int>(p, 4));
  });
  return 0;
}

extern "C" __attribute__((visibility("default"))) int mock_run_iree_b200_mxfp4_pack_scales(
    const unsigned grid[3], const unsigned block[3], void** p) {
  if (block[0] != 256) return 1;
  run_grid(grid, [&] {
    iree_b200_mxfp4_pack_scales(arg<const uint8_t*>(p, 0), arg<uint32_t*>(p, 1), arg<int>(p, 2),
                                arg<int>(p, 3));
  });
  return 0;
}
