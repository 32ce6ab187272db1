// iree_b200_tma_recipe.h — self-describing TMA requirements of a dispatch kernel.
//
// IREE's CUDA dispatch ABI passes only buffer pointers and 32-bit push constants
// (reference: runtime/src/iree/hal/drivers/cuda/stream_command_buffer.c:532-587,
// samples/custom_dispatch/cuda/kernels/kernels.cu:7-34). Blackwell kernels are fed by
// TMA, which needs a 128-byte CUtensorMap per tensor, encoded on the host from the
// device pointer and shape. To keep the kernels self-contained hal.executable.objects,
// every TMA-fed kernel `foo` ships a `__device__ const iree_b200_tma_recipe_t foo__tma`
// global in the same cubin. The thin HAL layer reads it once at executable load
// (cuModuleGetGlobal, the same mechanism as the reference's
// iree_hal_cuda_native_executable_lookup_global_by_name, native_executable.c:579-649),
// and at dispatch time encodes the maps from (binding pointers, push constants) and
// appends them as trailing `__grid_constant__` kernel parameters after the ABI ones.
//
// Plain C, shared by device code (kernels/*.cu) and host code (hal/*.c).

#ifndef IREE_B200_TMA_RECIPE_H_
#define IREE_B200_TMA_RECIPE_H_

#include <stdint.h>

#define IREE_B200_TMA_RECIPE_MAGIC 0x414D5442u /* "BTMA" */
#define IREE_B200_TMA_MAX_MAPS 6
#define IREE_B200_TMA_MAX_RANK 4

// Element types (values match CUtensorMapDataType).
enum {
  IREE_B200_TMA_U8 = 0,
  IREE_B200_TMA_U16 = 1,
  IREE_B200_TMA_U32 = 2,
  IREE_B200_TMA_I32 = 3,
  IREE_B200_TMA_F16 = 6,
  IREE_B200_TMA_F32 = 7,
  IREE_B200_TMA_BF16 = 9,
};

// Swizzle modes (values match CUtensorMapSwizzle).
enum {
  IREE_B200_TMA_SWIZZLE_NONE = 0,
  IREE_B200_TMA_SWIZZLE_32B = 1,
  IREE_B200_TMA_SWIZZLE_64B = 2,
  IREE_B200_TMA_SWIZZLE_128B = 3,
};

// A dimension extent is either a literal or "push constant #i".
#define IREE_B200_TMA_DIM_CONST(i) (0x80000000u | (uint32_t)(i))
#define IREE_B200_TMA_DIM_LITERAL(v) ((uint32_t)(v))

// Dense position of map dimension d (0 = innermost), 4 bits per dimension.
#define IREE_B200_TMA_ORDER(o0, o1, o2, o3) \
  ((uint32_t)(o0) | ((uint32_t)(o1) << 4) | ((uint32_t)(o2) << 8) | ((uint32_t)(o3) << 12))

typedef struct iree_b200_tma_map_t {
  uint32_t binding;    // index of the binding whose device pointer is the base
  uint32_t elem_type;  // IREE_B200_TMA_*
  uint32_t elem_size;  // bytes
  uint32_t rank;       // 1..4
  // Extents of the MAP's dimensions (a literal or "push constant #i"). The tensor is dense;
  // with |order| == 0 its layout is the map's own dimension order, innermost first:
  // stride[d] = elem_size * prod(dim[0..d)). A non-zero |order| (IREE_B200_TMA_ORDER) gives,
  // per map dimension, its position in the dense layout, so a map can walk a tiled tensor in
  // another order than it is stored: stride[d] = elem_size * prod(dim[e] : order[e] < order[d]).
  // Dimension 0 must be the dense-innermost one (order[0] == 0).
  uint32_t dim[IREE_B200_TMA_MAX_RANK];
  // Tile extents per map dimension (a literal or "push constant #i").
  uint32_t box[IREE_B200_TMA_MAX_RANK];
  uint32_t swizzle;                      // IREE_B200_TMA_SWIZZLE_* (a literal or "push constant #i")
  uint32_t order;                        // 0 = identity
} iree_b200_tma_map_t;

typedef struct iree_b200_tma_recipe_t {
  uint32_t magic;      // IREE_B200_TMA_RECIPE_MAGIC
  uint32_t map_count;  // number of trailing CUtensorMap parameters
  iree_b200_tma_map_t maps[IREE_B200_TMA_MAX_MAPS];
} iree_b200_tma_recipe_t;

#endif  // IREE_B200_TMA_RECIPE_H_
