// iree_b200_tma_hook.h — what a CUDA HAL driver adds to launch the TMA-fed sm_100a kernels.
//
// The kernel ABI passes pointers and i32 push constants only
// (runtime/src/iree/hal/drivers/cuda/stream_command_buffer.c:532-587). A TMA-fed kernel
// `foo` carries a `__device__ const iree_b200_tma_recipe_t foo__tma` global in its cubin
// (include/iree_b200_tma_recipe.h). This header is the driver side of that contract, as two
// calls a driver makes with its own dynamically resolved CUDA entry points:
//   * iree_b200_tma_hook_load    — once per export, right after cuModuleGetFunction
//                                  (native_executable.c:413): reads the recipe, if any;
//   * iree_b200_tma_hook_encode  — per dispatch, after bindings and constants are packed
//                                  (stream_command_buffer.c:587, graph_command_buffer.c:814):
//                                  encodes the CUtensorMaps that become the trailing
//                                  `__grid_constant__` kernel parameters.
// integration/cuda_driver_tma.patch wires them into the reference driver; the thin layer's
// own launch path (iree_b200/csrc/hal/executable.cc: encode_tensor_map) is the same logic.
//
// Plain C, header-only, no dependency on either code base beyond <cuda.h>.

#ifndef IREE_B200_TMA_HOOK_H_
#define IREE_B200_TMA_HOOK_H_

#include <cuda.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "iree_b200_tma_recipe.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef CUresult (*iree_b200_pfn_cuModuleGetGlobal)(CUdeviceptr*, size_t*, CUmodule, const char*);
typedef CUresult (*iree_b200_pfn_cuMemcpyDtoH)(void*, CUdeviceptr, size_t);
typedef CUresult (*iree_b200_pfn_cuTensorMapEncodeTiled)(
    CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
    const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// Reads `<kernel_name>__tma` from |module| into |out_recipe|. A kernel without a recipe (or
// with a malformed one) leaves map_count = 0: it launches exactly as before.
static inline void iree_b200_tma_hook_load(iree_b200_pfn_cuModuleGetGlobal get_global,
                                           iree_b200_pfn_cuMemcpyDtoH memcpy_dtoh,
                                           CUmodule module, const char* kernel_name,
                                           iree_b200_tma_recipe_t* out_recipe) {
  memset(out_recipe, 0, sizeof(*out_recipe));
  char symbol[256];
  if (snprintf(symbol, sizeof(symbol), "%s__tma", kernel_name) >= (int)sizeof(symbol)) return;
  CUdeviceptr ptr = 0;
  size_t size = 0;
  if (get_global(&ptr, &size, module, symbol) != CUDA_SUCCESS ||
      size != sizeof(iree_b200_tma_recipe_t)) {
    return;
  }
  iree_b200_tma_recipe_t recipe;
  if (memcpy_dtoh(&recipe, ptr, sizeof(recipe)) != CUDA_SUCCESS ||
      recipe.magic != IREE_B200_TMA_RECIPE_MAGIC || recipe.map_count > IREE_B200_TMA_MAX_MAPS) {
    return;
  }
  *out_recipe = recipe;
}

// Encodes map |i| of |recipe| from the dispatch's binding pointers and push constants.
// Returns CUDA_ERROR_INVALID_VALUE for a dispatch that breaks the TMA guards (which the
// compiler-side matcher excludes): unaligned base or pitch, zero extent, bad index.
static inline CUresult iree_b200_tma_hook_encode(iree_b200_pfn_cuTensorMapEncodeTiled encode_tiled,
                                                 const iree_b200_tma_recipe_t* recipe, uint32_t i,
                                                 const CUdeviceptr* binding_ptrs,
                                                 uint32_t binding_count, const uint32_t* constants,
                                                 uint32_t constant_count, CUtensorMap* out_map) {
  const iree_b200_tma_map_t* map = &recipe->maps[i];
  if (!encode_tiled) return CUDA_ERROR_NOT_SUPPORTED;  // driver older than CUDA 12
  if (map->rank < 1 || map->rank > IREE_B200_TMA_MAX_RANK || map->binding >= binding_count) {
    return CUDA_ERROR_INVALID_VALUE;
  }
  cuuint64_t dims[IREE_B200_TMA_MAX_RANK];
  cuuint64_t strides[IREE_B200_TMA_MAX_RANK];
  cuuint32_t box[IREE_B200_TMA_MAX_RANK];
  cuuint32_t elem_strides[IREE_B200_TMA_MAX_RANK];
  uint32_t order[IREE_B200_TMA_MAX_RANK];
  for (uint32_t d = 0; d < map->rank; ++d) {
    // Extents and box sizes are literals or "push constant #i".
    uint64_t extent = map->dim[d], tile = map->box[d];
    if (extent & 0x80000000u) {
      const uint32_t index = (uint32_t)(extent & 0x7fffffffu);
      if (index >= constant_count) return CUDA_ERROR_INVALID_VALUE;
      extent = constants[index];
    }
    if (tile & 0x80000000u) {
      const uint32_t index = (uint32_t)(tile & 0x7fffffffu);
      if (index >= constant_count) return CUDA_ERROR_INVALID_VALUE;
      tile = constants[index];
    }
    if (extent == 0 || tile == 0 || tile > 256) return CUDA_ERROR_INVALID_VALUE;
    dims[d] = extent;
    box[d] = (cuuint32_t)tile;
    elem_strides[d] = 1;
    order[d] = map->order ? (map->order >> (4 * d)) & 0xfu : d;
  }
  if (order[0] != 0) return CUDA_ERROR_INVALID_VALUE;
  for (uint32_t d = 1; d < map->rank; ++d) {
    // Dense tensor walked in the map's dimension order: the stride of a dimension is the
    // product of the extents stored inside it.
    uint64_t stride = map->elem_size;
    for (uint32_t e = 0; e < map->rank; ++e) {
      if (order[e] < order[d]) stride *= dims[e];
    }
    if (stride % 16 != 0) return CUDA_ERROR_INVALID_VALUE;
    strides[d - 1] = stride;  // globalStrides[j] = byte stride of dim j + 1
  }
  const CUdeviceptr base = binding_ptrs[map->binding];
  if (base % 16 != 0) return CUDA_ERROR_INVALID_VALUE;
  uint32_t swizzle = map->swizzle;
  if (swizzle & 0x80000000u) {
    const uint32_t index = swizzle & 0x7fffffffu;
    if (index >= constant_count) return CUDA_ERROR_INVALID_VALUE;
    swizzle = constants[index];
  }
  if (swizzle > IREE_B200_TMA_SWIZZLE_128B) return CUDA_ERROR_INVALID_VALUE;
  // The hardware faults on a swizzled box whose inner row is narrower than the swizzle span.
  if (swizzle && box[0] * map->elem_size < (16u << swizzle)) return CUDA_ERROR_INVALID_VALUE;
  return encode_tiled(out_map, (CUtensorMapDataType)map->elem_type, map->rank, (void*)base, dims,
                      strides, box, elem_strides, CU_TENSOR_MAP_INTERLEAVE_NONE,
                      (CUtensorMapSwizzle)swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
}

#ifdef __cplusplus
}  // extern "C"
#endif

#endif  // IREE_B200_TMA_HOOK_H_
