// manifest_tool.cc — prints the export table of the builtin executable as JSON.
//
// Plays the role of the hal.executable.export ops a compiler would carry: name, ordinal,
// workgroup_size, workgroup_local_memory, #hal.pipeline.layout (constants, bindings with
// their ReadOnly flag) — the inputs of CUDATarget.cpp:650-691. Built and run on the host
// by iree_b200/build.py; the geometry comes from the same constexpr definitions the
// kernels are compiled from.

#include <cstdio>
#include <initializer_list>

#include "../kernels/attention_config.h"
#include "../kernels/collective_config.h"
#include "../kernels/gemm_config.h"

using namespace b200;

static bool first = true;

static void emit(const char* name, const char* module, int threads, int smem, int constants,
                 const char* binding_flags) {
  printf("%s\n    {\"name\": \"%s\", \"module\": \"%s\", \"block_dims\": [%d, 1, 1], "
         "\"shared_memory\": %d, \"constants\": %d, \"binding_flags\": [%s]}",
         first ? "" : ",", name, module, threads, smem, constants, binding_flags);
  first = false;
}

int main() {
  printf("{\n  \"exports\": [");
  // READ_ONLY = 1 (cuda_executable_def.fbs:23-26).
#define X(NAME, CLUSTER, ...)                                                              \
  emit(#NAME, "gemm_kernels", GemmConfig<__VA_ARGS__>::kThreads,                           \
       GemmConfig<__VA_ARGS__>::kSmemBytes, 5, "1, 1, 0");
#include "../kernels/gemm_exports.inc"
#undef X
  // Same kernels, tensor maps built on the device: the stock custom-dispatch ABI, no recipe.
#define X(NAME, CLUSTER, ...)                                                              \
  emit(#NAME "_dtm", "gemm_dtm_kernels", GemmConfig<__VA_ARGS__>::kThreads,                \
       GemmConfig<__VA_ARGS__>::kSmemBytes, 5, "1, 1, 0");
#include "../kernels/gemm_exports.inc"
#undef X
#define X(NAME, CLUSTER, ...)                                                              \
  emit(#NAME "_dtm", "gemm_bias_dtm_kernels", GemmConfig<__VA_ARGS__>::kThreads,           \
       GemmConfig<__VA_ARGS__>::kSmemBytes, 5, "1, 1, 1, 0");
#include "../kernels/gemm_bias_exports.inc"
#undef X
#define X(NAME, CLUSTER, ...)                                                              \
  emit(#NAME, "gemm_bias_kernels", GemmConfig<__VA_ARGS__>::kThreads,                      \
       GemmConfig<__VA_ARGS__>::kSmemBytes, 5, "1, 1, 1, 0");
#include "../kernels/gemm_bias_exports.inc"
#undef X
  emit("iree_b200_matmul_generic_bf16_f32", "generic_kernels", 256, 0, 5, "1, 1, 0");
  emit("iree_b200_matmul_generic_f16_f32", "generic_kernels", 256, 0, 5, "1, 1, 0");
  emit("iree_b200_matmul_generic_i8_i32", "generic_kernels", 256, 0, 5, "1, 1, 0");
  emit("iree_b200_matmul_generic_f8e4m3fn_f32", "generic_kernels", 256, 0, 5, "1, 1, 0");
  emit("iree_b200_matmul_generic_f8e5m2_f32", "generic_kernels", 256, 0, 5, "1, 1, 0");
  emit("iree_b200_matmul_generic_bias_bf16_f32", "generic_kernels", 256, 0, 5, "1, 1, 1, 0");
  emit("iree_b200_matmul_generic_bias_f16_f32", "generic_kernels", 256, 0, 5, "1, 1, 1, 0");
  emit("iree_b200_matmul_generic_bias_i8_i32", "generic_kernels", 256, 0, 5, "1, 1, 1, 0");
#define X(NAME, THREADS, SMEM, CONSTANTS, BINDINGS) \
  emit(#NAME, "attention_kernels", THREADS, SMEM, CONSTANTS, BINDINGS);
#include "../kernels/attention_exports.inc"
#undef X
  // Kernels awaiting their first GPU run ship in a second executable that is loaded on demand
  // (build.py EXPERIMENTAL_UNITS; mx_kernels.cu): lhs, lhs_scales, rhs, rhs_scales, out.
  emit("iree_b200_matmul_generic_mxfp4_f32", "mx_kernels", 256, 0, 5, "1, 1, 1, 1, 0");
  // The block-scaled tcgen05 form: constants M, N, K, flags, K / 2.
  emit("iree_b200_matmul_mxfp4_f32", "mx_kernels", MxGemmConfig<2>::kThreads,
       MxGemmConfig<2>::kSmemBytes, 5, "1, 1, 1, 1, 0");
  emit("iree_b200_matmul_mxfp4_f32_1cta", "mx_kernels", MxGemmConfig<1>::kThreads,
       MxGemmConfig<1>::kSmemBytes, 5, "1, 1, 1, 1, 0");
  // ... with the scales already in the instruction's atom order: + K/128, ceil(M/128), ceil(N/128).
  emit("iree_b200_matmul_mxfp4_packed_f32", "mx_kernels", MxGemmConfig<2>::kThreads,
       MxGemmConfig<2>::kSmemBytes, 8, "1, 1, 1, 1, 0");
  emit("iree_b200_matmul_mxfp4_packed_f32_1cta", "mx_kernels", MxGemmConfig<1>::kThreads,
       MxGemmConfig<1>::kSmemBytes, 8, "1, 1, 1, 1, 0");
  // row-major scales -> atoms: scales, atoms; constants rows, K.
  emit("iree_b200_mxfp4_pack_scales", "mx_kernels", 256, 0, 2, "1, 0");
  // MXFP4 -> bf16 operand expansion: packed, scales, out; constants rows, K.
  emit("iree_b200_matmul_mxfp4_dequant_bf16", "mx_kernels", 256, 0, 2, "1, 1, 0");
  // linalg.mmt4d on data-tiled operands (mmt4d_kernels.cu): constants M1 N1 K1 M0 N0 K0 flags.
  emit("iree_b200_mmt4d_generic_bf16_f32", "mmt4d_kernels", 256, 0, 7, "1, 1, 0");
  emit("iree_b200_mmt4d_generic_f16_f32", "mmt4d_kernels", 256, 0, 7, "1, 1, 0");
  emit("iree_b200_mmt4d_generic_i8_i32", "mmt4d_kernels", 256, 0, 7, "1, 1, 0");
  emit("iree_b200_mmt4d_generic_f32_f32", "mmt4d_kernels", 256, 0, 7, "1, 1, 0");
  // ... on the tensor cores, tiles read and written directly (K0 rows of 16 bytes):
  // + the seven map extents of make_gemm_tiled_recipe.
  {
    using C2 = GemmConfig<2, KIND_F16, FMT_BF16, true, 6, false, false, true>;
    using C1 = GemmConfig<1, KIND_F16, FMT_BF16, true, 4, false, false, true>;
    for (const char* t : {"bf16_f32", "f16_f32", "i8_i32"}) {
      char name[96];
      snprintf(name, sizeof(name), "iree_b200_mmt4d_%s", t);
      emit(name, "mmt4d_kernels", C2::kThreads, C2::kSmemBytes, 14, "1, 1, 0");
      snprintf(name, sizeof(name), "iree_b200_mmt4d_%s_1cta", t);
      emit(name, "mmt4d_kernels", C1::kThreads, C1::kSmemBytes, 14, "1, 1, 0");
    }
  }
  // tiled <-> row-major relayout: src, dst; constants R1 C1 R0 C0 to_tiled.
  emit("iree_b200_mmt4d_relayout_b8", "mmt4d_kernels", 256, 0, 5, "1, 0");
  emit("iree_b200_mmt4d_relayout_b16", "mmt4d_kernels", 256, 0, 5, "1, 0");
  emit("iree_b200_mmt4d_relayout_b32", "mmt4d_kernels", 256, 0, 5, "1, 0");
  // Channel-launched kernels (not dispatch-ABI exports; see collective_kernels.cu).
  for (const char* type : {"f16", "bf16", "f32"}) {
    for (const char* path : {"nvls", "p2p"}) {
      char name[96];
      snprintf(name, sizeof(name), "iree_b200_all_reduce_sum_%s_%s", type, path);
      emit(name, "collective_kernels", CollectiveConfig::kThreads, 0, 0, "1");
    }
  }
  printf("\n  ]\n}\n");
  return 0;
}
