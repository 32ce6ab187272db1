// ops.cc — export selection ("matcher guards + count region") and dispatch recording for
// the dense-contraction ops. See include/iree_b200_ops.h.

#include <algorithm>
#include <mutex>
#include <vector>

#include "../hal/internal.h"
#include "../kernels/attention_config.h"
#include "../kernels/gemm_config.h"
#include "iree_b200_ops.h"

using namespace b200hal;

extern "C" {
// Provided by builtin_blob.S (.incbin of the packed CDA1 executable).
extern const uint8_t iree_b200_builtin_cda1_start[];
extern const uint8_t iree_b200_builtin_cda1_end[];
// The executable of kernels that have not run on a GPU yet (build.py EXPERIMENTAL_UNITS).
extern const uint8_t iree_b200_builtin_pending_cda1_start[];
extern const uint8_t iree_b200_builtin_pending_cda1_end[];
}

namespace {

enum MatmulFamily { kFamily2Cta = 0, kFamily1Cta = 1, kFamilyGeneric = 2, kFamilyCount = 3 };

// Indexed by IREE_B200_ELEMENT_TYPE_*; f32 operands have no matmul export.
constexpr int kTypeCount = 6;
const char* const kTypeSuffix[kTypeCount] = {"i8_i32",  "f16_f32",      "bf16_f32",
                                             nullptr,   "f8e4m3fn_f32", "f8e5m2_f32"};
const char* const kTruncSuffix[kTypeCount] = {nullptr, "f16_f16", "bf16_bf16",
                                              nullptr, nullptr,   nullptr};

bool matmul_type_ok(int32_t t) { return t >= 0 && t < kTypeCount && kTypeSuffix[t]; }
uint64_t matmul_elem_size(int32_t t) {
  return (t == IREE_B200_ELEMENT_TYPE_F16 || t == IREE_B200_ELEMENT_TYPE_BF16) ? 2 : 1;
}

// The exp2 split the prefill kernel uses unless a flag says otherwise.
constexpr int kDefaultPrefillVariant = 0;  // kAttnPrefill

enum AttentionExport {
  kAttnPrefill = 0,
  kAttnDecodeSplit = 1,
  kAttnDecodeMerge = 2,
  kAttnGeneric = 3,
  kAttnDecodeCluster = 4,
  kAttnPrefillPoly25 = 5,
  kAttnPrefillPoly50 = 6,
  kAttnPrefillSplit = 7,
  kAttnPrefillMerge = 8,
  kAttnPrefillMask = 9,
  kAttnPrefillDeviceMaps = 10
};

}  // namespace

struct iree_b200_ops_t {
  iree_b200_device_t* device = nullptr;
  iree_b200_executable_cache_t* cache = nullptr;
  iree_b200_executable_t* executable = nullptr;
  // Loaded by the first MX call only (iree_b200_ops_matmul_mxfp4_*).
  iree_b200_executable_t* pending_executable = nullptr;
  int32_t mxfp4_generic_export = -1;
  int32_t mxfp4_dequant_export = -1;
  int32_t mxfp4_block_scaled_export = -1;
  int32_t mxfp4_block_scaled_1cta_export = -1;
  int32_t mxfp4_packed_export = -1, mxfp4_packed_1cta_export = -1, mxfp4_pack_scales_export = -1;
  int32_t mmt4d_generic_exports[4] = {-1, -1, -1, -1};  // I8, F16, BF16, F32
  int32_t mmt4d_direct_exports[3][2] = {{-1, -1}, {-1, -1}, {-1, -1}};  // [I8, F16, BF16][2-CTA, 1-CTA]
  int32_t mmt4d_relayout_exports[3] = {-1, -1, -1};  // 1-, 2-, 4-byte elements
  std::mutex pending_mutex;
  std::atomic<uint64_t> dispatch_count{0};
  int32_t reserved_sms = 0;
  int32_t sm_count = 0;  // device->sm_count, or the caller's for an offline (selection-only) object
  // [element_type][transpose_rhs][family] → export ordinal (-1 if absent)
  int32_t matmul_exports[6][2][kFamilyCount];
  // The `_dtm` twins (tensor maps built on the device): [0] plain, [1] truncating,
  // [2] bias, [3] bias + truncating.
  int32_t matmul_dtm_exports[4][6][2][kFamilyCount];
  int32_t matmul_trunc_exports[6][2][kFamilyCount];  // truncating epilogue (f16 / bf16)
  int32_t matmul_bias_exports[6][2][kFamilyCount];        // bias / activation epilogue
  int32_t matmul_bias_trunc_exports[6][2][kFamilyCount];  // ... then truncation
  int32_t attention_exports[4][12];
  // Host-buffer calls: staging arena, transfer streams and events (grow-only, reused).
  std::mutex host_mutex;
  uint64_t pipeline_min_bytes = 16ull << 20;
  int32_t pipeline_block_m = 2048, pipeline_block_n = 2048;
  iree_b200_buffer_t* arena = nullptr;
  CUstream upload_stream = nullptr, download_stream = nullptr;
  std::vector<CUevent> events;
};

namespace {

int32_t lookup(iree_b200_executable_t* exe, const std::string& name) {
  int32_t ordinal = -1;
  if (iree_b200_executable_lookup_export_by_name(exe, name.c_str(), &ordinal) !=
      IREE_B200_STATUS_OK) {
    return -1;
  }
  return ordinal;
}

int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }

}  // namespace

extern "C" {

const void* iree_b200_ops_builtin_executable_data(size_t* out_size) {
  if (out_size) *out_size = (size_t)(iree_b200_builtin_cda1_end - iree_b200_builtin_cda1_start);
  return iree_b200_builtin_cda1_start;
}

// Export ordinals of every kernel family, looked up by name in ops->executable.
static void populate_export_tables(iree_b200_ops_t* ops) {
  for (int t = 0; t < kTypeCount; ++t) {
    for (int nt = 0; nt < 2; ++nt) {
      for (int f = 0; f < kFamilyCount; ++f) {
        ops->matmul_exports[t][nt][f] = ops->matmul_trunc_exports[t][nt][f] = -1;
        for (int v = 0; v < 4; ++v) ops->matmul_dtm_exports[v][t][nt][f] = -1;
        ops->matmul_bias_exports[t][nt][f] = ops->matmul_bias_trunc_exports[t][nt][f] = -1;
      }
      if (!kTypeSuffix[t]) continue;
      const std::string base = std::string("iree_b200_matmul_") + (nt ? "nt_" : "") + kTypeSuffix[t];
      ops->matmul_exports[t][nt][kFamily2Cta] = lookup(ops->executable, base);
      ops->matmul_exports[t][nt][kFamily1Cta] = lookup(ops->executable, base + "_1cta");
      ops->matmul_dtm_exports[0][t][nt][kFamily2Cta] = lookup(ops->executable, base + "_dtm");
      ops->matmul_dtm_exports[0][t][nt][kFamily1Cta] = lookup(ops->executable, base + "_1cta_dtm");
      ops->matmul_exports[t][nt][kFamilyGeneric] =
          lookup(ops->executable, std::string("iree_b200_matmul_generic_") + kTypeSuffix[t]);
      if (t <= IREE_B200_ELEMENT_TYPE_BF16) {  // i8, f16, bf16
        const std::string bbase =
            std::string("iree_b200_matmul_bias_") + (nt ? "nt_" : "") + kTypeSuffix[t];
        ops->matmul_bias_exports[t][nt][kFamily2Cta] = lookup(ops->executable, bbase);
        ops->matmul_bias_exports[t][nt][kFamily1Cta] = lookup(ops->executable, bbase + "_1cta");
        ops->matmul_bias_exports[t][nt][kFamilyGeneric] = lookup(
            ops->executable, std::string("iree_b200_matmul_generic_bias_") + kTypeSuffix[t]);
        ops->matmul_dtm_exports[2][t][nt][kFamily2Cta] = lookup(ops->executable, bbase + "_dtm");
        ops->matmul_dtm_exports[2][t][nt][kFamily1Cta] =
            lookup(ops->executable, bbase + "_1cta_dtm");
      }
      if (!kTruncSuffix[t]) continue;
      const std::string btbase =
          std::string("iree_b200_matmul_bias_") + (nt ? "nt_" : "") + kTruncSuffix[t];
      ops->matmul_bias_trunc_exports[t][nt][kFamily2Cta] = lookup(ops->executable, btbase);
      ops->matmul_bias_trunc_exports[t][nt][kFamily1Cta] =
          lookup(ops->executable, btbase + "_1cta");
      ops->matmul_bias_trunc_exports[t][nt][kFamilyGeneric] =
          ops->matmul_bias_exports[t][nt][kFamilyGeneric];
      ops->matmul_dtm_exports[3][t][nt][kFamily2Cta] = lookup(ops->executable, btbase + "_dtm");
      ops->matmul_dtm_exports[3][t][nt][kFamily1Cta] =
          lookup(ops->executable, btbase + "_1cta_dtm");
      const std::string tbase =
          std::string("iree_b200_matmul_") + (nt ? "nt_" : "") + kTruncSuffix[t];
      ops->matmul_trunc_exports[t][nt][kFamily2Cta] = lookup(ops->executable, tbase);
      ops->matmul_trunc_exports[t][nt][kFamily1Cta] = lookup(ops->executable, tbase + "_1cta");
      ops->matmul_trunc_exports[t][nt][kFamilyGeneric] = ops->matmul_exports[t][nt][kFamilyGeneric];
      ops->matmul_dtm_exports[1][t][nt][kFamily2Cta] = lookup(ops->executable, tbase + "_dtm");
      ops->matmul_dtm_exports[1][t][nt][kFamily1Cta] = lookup(ops->executable, tbase + "_1cta_dtm");
    }
  }
  for (int t = 0; t < 4; ++t) {
    for (int e = 0; e < 12; ++e) ops->attention_exports[t][e] = -1;
  }
  ops->attention_exports[IREE_B200_ELEMENT_TYPE_F32][kAttnGeneric] =
      lookup(ops->executable, "iree_b200_attention_generic_f32");
  for (int t = 1; t < 3; ++t) {
    const std::string suffix = t == IREE_B200_ELEMENT_TYPE_F16 ? "f16" : "bf16";
    ops->attention_exports[t][kAttnPrefill] =
        lookup(ops->executable, "iree_b200_attention_prefill_" + suffix);
    ops->attention_exports[t][kAttnPrefillPoly25] =
        lookup(ops->executable, "iree_b200_attention_prefill_" + suffix + "_p25");
    ops->attention_exports[t][kAttnPrefillPoly50] =
        lookup(ops->executable, "iree_b200_attention_prefill_" + suffix + "_p50");
    ops->attention_exports[t][kAttnPrefillSplit] =
        lookup(ops->executable, "iree_b200_attention_prefill_split_" + suffix);
    ops->attention_exports[t][kAttnPrefillMerge] =
        lookup(ops->executable, "iree_b200_attention_prefill_merge_" + suffix);
    ops->attention_exports[t][kAttnPrefillMask] =
        lookup(ops->executable, "iree_b200_attention_prefill_mask_" + suffix);
    ops->attention_exports[t][kAttnPrefillDeviceMaps] =
        lookup(ops->executable, "iree_b200_attention_prefill_" + suffix + "_dtm");
    ops->attention_exports[t][kAttnDecodeSplit] =
        lookup(ops->executable, "iree_b200_attention_decode_split_" + suffix);
    ops->attention_exports[t][kAttnDecodeMerge] =
        lookup(ops->executable, "iree_b200_attention_decode_merge_" + suffix);
    ops->attention_exports[t][kAttnDecodeCluster] =
        lookup(ops->executable, "iree_b200_attention_decode_cluster_" + suffix);
    ops->attention_exports[t][kAttnGeneric] =
        lookup(ops->executable, "iree_b200_attention_generic_" + suffix);
  }
}

iree_b200_status_t iree_b200_ops_create_from_data(iree_b200_device_t* device, const void* data,
                                                  size_t size, iree_b200_ops_t** out_ops) {
  B200_REQUIRE_ARG(device && data && out_ops);
  *out_ops = nullptr;
  if (device->cc_major != 10) {
    return make_status(IREE_B200_STATUS_INCOMPATIBLE,
                       "the builtin kernels are sm_100a only; device %d is sm_%d%d", device->ordinal,
                       device->cc_major, device->cc_minor);
  }
  auto* ops = new iree_b200_ops_t();
  ops->device = device;
  iree_b200_device_retain(device);
  iree_b200_status_t status = iree_b200_executable_cache_create(device, "iree_b200_ops", &ops->cache);
  if (status == IREE_B200_STATUS_OK) {
    status = iree_b200_executable_cache_prepare_executable(ops->cache, "PTXE", data, size,
                                                           &ops->executable);
  }
  if (status != IREE_B200_STATUS_OK) {
    std::string message = iree_b200_status_message();
    iree_b200_ops_release(ops);
    return make_status(status, "%s", message.c_str());
  }
  ops->sm_count = device->sm_count;
  populate_export_tables(ops);
  *out_ops = ops;
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t iree_b200_ops_create_offline(const void* data, size_t size, int32_t sm_count,
                                                iree_b200_ops_t** out_ops) {
  B200_REQUIRE_ARG(out_ops);
  *out_ops = nullptr;
  if (sm_count < 2) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "sm_count %d is not a device", sm_count);
  }
  if (!data) data = iree_b200_ops_builtin_executable_data(&size);
  const uint8_t* fb = nullptr;
  size_t fb_length = 0;
  B200_RETURN_IF_ERROR(read_executable_header(data, size, /*unsafe_infer_size=*/false, &fb,
                                              &fb_length));
  ExecutableDefView view;
  B200_RETURN_IF_ERROR(parse_executable_def(fb, fb_length, &view));
  auto* ops = new iree_b200_ops_t();
  ops->sm_count = sm_count;
  // An executable with names and layouts but no device, no modules and no functions: good
  // for lookups and export_info, never dispatchable (there is no command buffer without a
  // device, and the record / host entry points refuse an offline object).
  ops->executable = new iree_b200_executable_t();
  ops->executable->exports.resize(view.exports.size());
  for (size_t i = 0; i < view.exports.size(); ++i) {
    const ExportDefView& e = view.exports[i];
    KernelParams& k = ops->executable->exports[i];
    k.name = e.kernel_name;
    for (int d = 0; d < 3; ++d) k.block_dims[d] = e.block_dims[d];
    k.block_shared_memory_size = e.block_shared_memory_size;
    k.constant_count = e.constant_count;
    k.binding_count = (uint32_t)e.binding_flags.size();
    for (size_t j = 0; j < e.binding_flags.size(); ++j) k.binding_flags[j] = e.binding_flags[j];
  }
  populate_export_tables(ops);
  *out_ops = ops;
  return IREE_B200_STATUS_OK;
}

// Every entry point that touches the GPU starts with this.
static iree_b200_status_t require_device(const iree_b200_ops_t* ops) {
  if (!ops) return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "ops is NULL");
  if (!ops->device) {
    return make_status(IREE_B200_STATUS_FAILED_PRECONDITION,
                       "this ops object was created offline: it selects kernels but cannot "
                       "record or run them (there is no CPU path)");
  }
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t iree_b200_ops_create(iree_b200_device_t* device, iree_b200_ops_t** out_ops) {
  size_t size = 0;
  const void* data = iree_b200_ops_builtin_executable_data(&size);
  return iree_b200_ops_create_from_data(device, data, size, out_ops);
}

void iree_b200_ops_release(iree_b200_ops_t* ops) {
  if (!ops) return;
  if (ops->device && ops->device->cu_context) {
    const CudaSymbols* cu = ops->device->cu;
    cu->cuCtxSetCurrent(ops->device->cu_context);
    for (CUstream s : {ops->upload_stream, ops->download_stream}) {
      if (!s) continue;
      cu->cuStreamSynchronize(s);
      cu->cuStreamDestroy(s);
    }
    for (CUevent e : ops->events) cu->cuEventDestroy(e);
  }
  iree_b200_buffer_release(ops->arena);
  iree_b200_executable_release(ops->executable);
  iree_b200_executable_release(ops->pending_executable);
  iree_b200_executable_cache_release(ops->cache);
  iree_b200_device_release(ops->device);
  delete ops;
}

iree_b200_executable_t* iree_b200_ops_executable(iree_b200_ops_t* ops) {
  return ops ? ops->executable : nullptr;
}

iree_b200_status_t iree_b200_ops_set_reserved_sms(iree_b200_ops_t* ops, int32_t reserved_sms) {
  B200_REQUIRE_ARG(ops);
  if (reserved_sms < 0 || reserved_sms >= ops->sm_count) {
    return make_status(IREE_B200_STATUS_OUT_OF_RANGE, "reserved_sms %d out of range [0, %d)",
                       reserved_sms, ops->sm_count);
  }
  ops->reserved_sms = reserved_sms;
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t iree_b200_ops_set_host_pipeline(iree_b200_ops_t* ops, uint64_t min_bytes,
                                                   int32_t block_m, int32_t block_n) {
  B200_REQUIRE_ARG(ops);
  if (block_m < 256 || block_n < 256 || block_m > (1 << 20) || block_n > (1 << 20)) {
    return make_status(IREE_B200_STATUS_OUT_OF_RANGE,
                       "host pipeline blocks must be within [256, 2^20] (got %d x %d)", block_m,
                       block_n);
  }
  std::lock_guard<std::mutex> lock(ops->host_mutex);
  ops->pipeline_min_bytes = min_bytes;
  ops->pipeline_block_m = (block_m + 255) / 256 * 256;
  ops->pipeline_block_n = (block_n + 255) / 256 * 256;
  return IREE_B200_STATUS_OK;
}

uint64_t iree_b200_ops_dispatch_count(const iree_b200_ops_t* ops) {
  return ops ? ops->dispatch_count.load() : 0;
}

iree_b200_status_t iree_b200_ops_select_matmul(const iree_b200_ops_t* ops, int32_t element_type,
                                               int64_t batch, int64_t M, int64_t N, int64_t K,
                                               uint32_t flags, uint32_t alignment,
                                               iree_b200_matmul_selection_t* out_selection) {
  B200_REQUIRE_ARG(ops && out_selection);
  memset(out_selection, 0, sizeof(*out_selection));
  if (!matmul_type_ok(element_type)) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "unsupported matmul element type %d",
                       element_type);
  }
  if (batch < 0 || M < 0 || N < 0 || K < 0 || batch > INT32_MAX || M > INT32_MAX ||
      N > INT32_MAX || K > INT32_MAX) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                       "matmul shape %lldx%lldx%lldx%lld out of range", (long long)batch,
                       (long long)M, (long long)N, (long long)K);
  }
  const bool transpose_rhs = (flags & IREE_B200_MATMUL_FLAG_TRANSPOSE_RHS) != 0;
  const bool truncate = (flags & IREE_B200_MATMUL_FLAG_TRUNCATE_RESULT) != 0;
  const int64_t es = (int64_t)matmul_elem_size(element_type);
  const int64_t out_es = truncate ? 2 : 4;
  if (truncate && (!kTruncSuffix[element_type] || (flags & IREE_B200_MATMUL_FLAG_ACCUMULATE))) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                       "TRUNCATE_RESULT applies to f16/bf16 without ACCUMULATE only");
  }
  const bool with_bias = (flags & IREE_B200_MATMUL_FLAG_BIAS) != 0;
  const bool relu = (flags & IREE_B200_MATMUL_FLAG_RELU) != 0;
  if (relu && !with_bias) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "FLAG_RELU needs FLAG_BIAS");
  }
  if (with_bias && element_type > IREE_B200_ELEMENT_TYPE_BF16) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                       "the bias / activation dispatches exist for i8, f16 and bf16 only");
  }
  out_selection->constants[0] = (uint32_t)M;
  out_selection->constants[1] = (uint32_t)N;
  out_selection->constants[2] = (uint32_t)K;
  out_selection->constants[3] = (uint32_t)batch;
  out_selection->constants[4] = (flags & IREE_B200_MATMUL_FLAG_ACCUMULATE) | (relu ? 2u : 0u);
  out_selection->export_ordinal = -1;
  if (batch == 0 || M == 0 || N == 0) return IREE_B200_STATUS_OK;  // empty result: no dispatch

  // Matcher guards of the TMA-fed kernels: 16-byte aligned bases and row pitches.
  const bool tma_ok = alignment >= 16 && K > 0 && (K * es) % 16 == 0 && (N * out_es) % 16 == 0 &&
                      (transpose_rhs || (N * es) % 16 == 0);
  const int sm_count = std::max(2, ops->sm_count - ops->reserved_sms);
  const int64_t tiles_2cta = ceil_div(M, 256) * ceil_div(N, 256) * batch;
  const int64_t tiles_1cta = ceil_div(M, 128) * ceil_div(N, 256) * batch;

  int family;
  if (flags & IREE_B200_MATMUL_FLAG_FORCE_GENERIC) {
    family = kFamilyGeneric;
  } else if (!tma_ok) {
    if (flags & (IREE_B200_MATMUL_FLAG_FORCE_1CTA | IREE_B200_MATMUL_FLAG_FORCE_2CTA)) {
      return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                         "shape/alignment does not satisfy the TMA guards of the forced kernel");
    }
    family = kFamilyGeneric;
  } else if (flags & IREE_B200_MATMUL_FLAG_FORCE_1CTA) {
    family = kFamily1Cta;
  } else if (flags & IREE_B200_MATMUL_FLAG_FORCE_2CTA) {
    family = kFamily2Cta;
  } else {
    // SM pairs only pay off once there are enough 256x256 tiles to occupy them.
    family = (M > 128 && tiles_2cta * 2 >= sm_count) ? kFamily2Cta : kFamily1Cta;
  }
  const bool device_maps = (flags & IREE_B200_MATMUL_FLAG_DEVICE_MAPS) != 0;
  if (device_maps && family == kFamilyGeneric) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                       "FLAG_DEVICE_MAPS: the shape must pass the TMA guards (the generic kernels "
                       "take no tensor maps to begin with)");
  }
  const auto& table = device_maps ? ops->matmul_dtm_exports[(with_bias ? 2 : 0) + (truncate ? 1 : 0)]
                      : with_bias ? (truncate ? ops->matmul_bias_trunc_exports : ops->matmul_bias_exports)
                                  : (truncate ? ops->matmul_trunc_exports : ops->matmul_exports);
  const int32_t ordinal = table[element_type][transpose_rhs ? 1 : 0][family];
  if (ordinal < 0) {
    return make_status(IREE_B200_STATUS_NOT_FOUND,
                       "the executable has no export for matmul type %d transpose_rhs %d family %d",
                       element_type, (int)transpose_rhs, family);
  }
  out_selection->export_ordinal = ordinal;
  switch (family) {
    case kFamily2Cta:
      // Persistent: one CTA pair per SM pair; the grid must be a multiple of the
      // compiled-in cluster size (the driver launches with cuLaunchKernel only).
      out_selection->workgroup_count[0] =
          (uint32_t)std::min<int64_t>(tiles_2cta * 2, sm_count & ~1);
      out_selection->workgroup_count[1] = 1;
      out_selection->workgroup_count[2] = 1;
      break;
    case kFamily1Cta:
      out_selection->workgroup_count[0] = (uint32_t)std::min<int64_t>(tiles_1cta, sm_count);
      out_selection->workgroup_count[1] = 1;
      out_selection->workgroup_count[2] = 1;
      break;
    default:
      if (batch > 65535 || ceil_div(M, 64) > 65535) {
        return make_status(IREE_B200_STATUS_OUT_OF_RANGE, "generic matmul grid out of range");
      }
      out_selection->workgroup_count[0] = (uint32_t)ceil_div(N, 64);
      out_selection->workgroup_count[1] = (uint32_t)ceil_div(M, 64);
      out_selection->workgroup_count[2] = (uint32_t)batch;
      // The generic kernels' own flag word: 1 accumulate, 2 transpose_rhs, 4 truncate, 8 relu.
      out_selection->constants[4] = (flags & IREE_B200_MATMUL_FLAG_ACCUMULATE) |
                                    (transpose_rhs ? 2u : 0u) | (truncate ? 4u : 0u) |
                                    (relu ? 8u : 0u);
      break;
  }
  return IREE_B200_STATUS_OK;
}

static uint32_t pointer_alignment(uint64_t p) {
  uint32_t a = 256;
  while (a > 1 && (p % a) != 0) a >>= 1;
  return a;
}

static iree_b200_status_t record_matmul_impl(iree_b200_ops_t* ops, iree_b200_command_buffer_t* cb,
                                             int32_t element_type, int64_t batch, int64_t M,
                                             int64_t N, int64_t K, uint32_t flags,
                                             const iree_b200_buffer_ref_t* refs, int ref_count) {
  B200_REQUIRE_ARG(ops && cb);
  B200_RETURN_IF_ERROR(require_device(ops));
  for (int i = 0; i < ref_count; ++i) B200_REQUIRE_ARG(refs[i].buffer);
  const bool with_bias = ref_count == 4;  // lhs, rhs, bias, out
  const uint64_t es = matmul_elem_size(element_type);
  // Binding extents must cover the operands (the HAL validation layer's job,
  // hal/command_buffer_validation.c).
  const uint64_t out_es = (flags & IREE_B200_MATMUL_FLAG_TRUNCATE_RESULT) ? 2 : 4;
  uint64_t need[4] = {(uint64_t)batch * M * K * es, (uint64_t)batch * K * N * es,
                      (uint64_t)batch * M * N * out_es, 0};
  if (with_bias) {
    need[3] = need[2];
    need[2] = (uint64_t)N * 4;
  }
  uint32_t alignment = 256;
  for (int i = 0; i < ref_count; ++i) {
    const uint64_t total = refs[i].buffer->byte_length;
    if (refs[i].offset > total) {
      return make_status(IREE_B200_STATUS_OUT_OF_RANGE, "matmul binding %d offset out of range", i);
    }
    const uint64_t length =
        refs[i].length == UINT64_MAX ? total - refs[i].offset : refs[i].length;
    if (length < need[i] || length > total - refs[i].offset) {
      return make_status(IREE_B200_STATUS_OUT_OF_RANGE,
                         "matmul binding %d holds %llu bytes but the shape needs %llu", i,
                         (unsigned long long)length, (unsigned long long)need[i]);
    }
    const uint64_t address = refs[i].buffer->device_ptr + refs[i].offset;
    if (with_bias && i == 2) {
      // The bias is read with scalar loads: element alignment is all it needs.
      if (address % 4) {
        return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "matmul bias must be 4-byte aligned");
      }
      continue;
    }
    alignment = std::min(alignment, pointer_alignment(address));
  }
  iree_b200_matmul_selection_t sel;
  B200_RETURN_IF_ERROR(
      iree_b200_ops_select_matmul(ops, element_type, batch, M, N, K, flags, alignment, &sel));
  if (sel.export_ordinal < 0) return IREE_B200_STATUS_OK;  // empty result
  B200_RETURN_IF_ERROR(iree_b200_command_buffer_dispatch(cb, ops->executable, sel.export_ordinal,
                                                         sel.workgroup_count, sel.constants, 5, refs,
                                                         ref_count, 0));
  ops->dispatch_count.fetch_add(1);
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t iree_b200_ops_record_matmul(iree_b200_ops_t* ops, iree_b200_command_buffer_t* cb,
                                               int32_t element_type, int64_t batch, int64_t M,
                                               int64_t N, int64_t K, uint32_t flags,
                                               iree_b200_buffer_ref_t lhs,
                                               iree_b200_buffer_ref_t rhs,
                                               iree_b200_buffer_ref_t out) {
  if (flags & (IREE_B200_MATMUL_FLAG_BIAS | IREE_B200_MATMUL_FLAG_RELU)) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                       "FLAG_BIAS dispatches are recorded with iree_b200_ops_record_matmul_bias");
  }
  const iree_b200_buffer_ref_t refs[3] = {lhs, rhs, out};
  return record_matmul_impl(ops, cb, element_type, batch, M, N, K, flags, refs, 3);
}

iree_b200_status_t iree_b200_ops_record_matmul_bias(
    iree_b200_ops_t* ops, iree_b200_command_buffer_t* cb, int32_t element_type, int64_t batch,
    int64_t M, int64_t N, int64_t K, uint32_t flags, iree_b200_buffer_ref_t lhs,
    iree_b200_buffer_ref_t rhs, iree_b200_buffer_ref_t bias, iree_b200_buffer_ref_t out) {
  const iree_b200_buffer_ref_t refs[4] = {lhs, rhs, bias, out};
  return record_matmul_impl(ops, cb, element_type, batch, M, N, K,
                            flags | IREE_B200_MATMUL_FLAG_BIAS, refs, 4);
}

namespace {

// Grow-only device staging for the host-buffer calls; caller holds ops->host_mutex.
iree_b200_status_t reserve_arena(iree_b200_ops_t* ops, uint64_t bytes) {
  if (ops->arena && ops->arena->byte_length >= bytes) return IREE_B200_STATUS_OK;
  iree_b200_buffer_release(ops->arena);
  ops->arena = nullptr;
  return iree_b200_allocator_allocate_buffer(ops->device, IREE_B200_MEMORY_TYPE_DEVICE_LOCAL,
                                             (bytes + (1ull << 20) - 1) >> 20 << 20, &ops->arena);
}

iree_b200_status_t ensure_transfer_streams(iree_b200_ops_t* ops) {
  const CudaSymbols* cu = ops->device->cu;
  B200_CU(cu, cuCtxSetCurrent(ops->device->cu_context));
  if (!ops->upload_stream) B200_CU(cu, cuStreamCreate(&ops->upload_stream, CU_STREAM_NON_BLOCKING));
  if (!ops->download_stream) {
    B200_CU(cu, cuStreamCreate(&ops->download_stream, CU_STREAM_NON_BLOCKING));
  }
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t event_at(iree_b200_ops_t* ops, size_t index, CUevent* out_event) {
  const CudaSymbols* cu = ops->device->cu;
  while (ops->events.size() <= index) {
    CUevent e = nullptr;
    B200_CU(cu, cuEventCreate(&e, CU_EVENT_DISABLE_TIMING));
    ops->events.push_back(e);
  }
  *out_event = ops->events[index];
  return IREE_B200_STATUS_OK;
}

// Strided host <-> dense device copy of |rows| x |width| bytes.
iree_b200_status_t copy_2d(const CudaSymbols* cu, bool to_device, CUdeviceptr device, void* host,
                           uint64_t host_pitch, uint64_t width, uint64_t rows, CUstream stream) {
  if (!width || !rows) return IREE_B200_STATUS_OK;
  if (host_pitch == width) {
    if (to_device) {
      B200_CU(cu, cuMemcpyHtoDAsync(device, host, width * rows, stream));
    } else {
      B200_CU(cu, cuMemcpyDtoHAsync(host, device, width * rows, stream));
    }
    return IREE_B200_STATUS_OK;
  }
  CUDA_MEMCPY2D p;
  memset(&p, 0, sizeof(p));
  if (to_device) {
    p.srcMemoryType = CU_MEMORYTYPE_HOST;
    p.srcHost = host;
    p.srcPitch = host_pitch;
    p.dstMemoryType = CU_MEMORYTYPE_DEVICE;
    p.dstDevice = device;
    p.dstPitch = width;
  } else {
    p.srcMemoryType = CU_MEMORYTYPE_DEVICE;
    p.srcDevice = device;
    p.srcPitch = width;
    p.dstMemoryType = CU_MEMORYTYPE_HOST;
    p.dstHost = host;
    p.dstPitch = host_pitch;
  }
  p.WidthInBytes = width;
  p.Height = rows;
  B200_CU(cu, cuMemcpy2DAsync(&p, stream));
  return IREE_B200_STATUS_OK;
}

// Upload / dispatch / download in sequence on the dispatch stream.
iree_b200_status_t matmul_host_serial(iree_b200_ops_t* ops, int32_t element_type, int64_t batch,
                                      int64_t M, int64_t N, int64_t K, uint32_t flags,
                                      const uint64_t bytes[3], const void* lhs, const void* rhs,
                                      void* out, std::vector<iree_b200_command_buffer_t*>* cbs,
                                      const void* bias = nullptr) {
  iree_b200_device_t* device = ops->device;
  const uint64_t out_es = (flags & IREE_B200_MATMUL_FLAG_TRUNCATE_RESULT) ? 2 : 4;
  const uint64_t off[3] = {0, (bytes[0] + 1023) / 1024 * 1024,
                           (bytes[0] + 1023) / 1024 * 1024 + (bytes[1] + 1023) / 1024 * 1024};
  const uint64_t off_bias = off[2] + (bytes[2] + 1023) / 1024 * 1024;
  const uint64_t bias_bytes = bias ? (uint64_t)N * 4 : 0;
  B200_RETURN_IF_ERROR(reserve_arena(ops, off_bias + bias_bytes + 1024));
  iree_b200_buffer_t* arena = ops->arena;
  if (bias_bytes) {
    B200_RETURN_IF_ERROR(iree_b200_device_transfer_h2d(device, bias, arena, off_bias, bias_bytes));
  }
  B200_RETURN_IF_ERROR(iree_b200_device_transfer_h2d(device, lhs, arena, off[0], bytes[0]));
  B200_RETURN_IF_ERROR(iree_b200_device_transfer_h2d(device, rhs, arena, off[1], bytes[1]));
  if (flags & IREE_B200_MATMUL_FLAG_ACCUMULATE) {
    B200_RETURN_IF_ERROR(iree_b200_device_transfer_h2d(device, out, arena, off[2], bytes[2]));
  }
  iree_b200_command_buffer_t* cb = nullptr;
  B200_RETURN_IF_ERROR(iree_b200_command_buffer_create(device, &cb));
  cbs->push_back(cb);
  B200_RETURN_IF_ERROR(iree_b200_command_buffer_begin(cb));
  if (bias) {
    // K == 0 included: the generic kernel then stores act((out +) bias).
    B200_RETURN_IF_ERROR(iree_b200_ops_record_matmul_bias(
        ops, cb, element_type, batch, M, N, K, flags,
        iree_b200_buffer_ref_t{arena, off[0], bytes[0]},
        iree_b200_buffer_ref_t{arena, off[1], bytes[1]},
        iree_b200_buffer_ref_t{arena, off_bias, bias_bytes},
        iree_b200_buffer_ref_t{arena, off[2], bytes[2]}));
  } else if (K == 0 && !(flags & IREE_B200_MATMUL_FLAG_ACCUMULATE) && bytes[2]) {
    // Empty reduction: the linalg.fill 0 that seeds the accumulator is the result.
    const uint32_t zero = 0;
    B200_RETURN_IF_ERROR(iree_b200_command_buffer_fill_buffer(
        cb, iree_b200_buffer_ref_t{arena, off[2], bytes[2]}, &zero, out_es == 2 ? 2 : 4));
  } else if (K > 0) {
    B200_RETURN_IF_ERROR(iree_b200_ops_record_matmul(
        ops, cb, element_type, batch, M, N, K, flags, iree_b200_buffer_ref_t{arena, off[0], bytes[0]},
        iree_b200_buffer_ref_t{arena, off[1], bytes[1]},
        iree_b200_buffer_ref_t{arena, off[2], bytes[2]}));
  }
  B200_RETURN_IF_ERROR(iree_b200_command_buffer_end(cb));
  B200_RETURN_IF_ERROR(
      iree_b200_device_queue_execute(device, 0, nullptr, nullptr, 0, nullptr, nullptr, cb));
  B200_RETURN_IF_ERROR(iree_b200_device_transfer_d2h(device, arena, off[2], out, bytes[2]));
  return iree_b200_device_wait_idle(device);
}

// The pipelined form (include/iree_b200_ops.h, iree_b200_ops_set_host_pipeline). Device
// staging is dense per block so every block is an ordinary dense dispatch:
//   A   [batch, M, K] as on the host; row panel (b, i) is contiguous
//   B   per batch, column panels [K, nb_j] packed one after the other (2-D upload);
//       with TRANSPOSE_RHS a panel is rows of [N, K] and already contiguous
//   C   per batch and row panel, blocks [mb_i, nb_j] packed one after the other (2-D download)
// Panel boundaries of one result dimension for the pipelined host call: whole blocks, then
// the remainder. (A half-height first panel, to shorten the fill, measured 3 % slower at
// 8192^3 — profiles/host_pipeline_sweep_r1.json — so the panels stay uniform.)
std::vector<int64_t> panel_edges(int64_t extent, int64_t block) {
  std::vector<int64_t> edges{0};
  while (edges.back() < extent) edges.push_back(std::min(extent, edges.back() + block));
  return edges;
}

iree_b200_status_t matmul_host_pipelined(iree_b200_ops_t* ops, int32_t element_type,
                                         int64_t batch, int64_t M, int64_t N, int64_t K,
                                         uint32_t flags, const uint64_t bytes[3], const void* lhs,
                                         const void* rhs, void* out,
                                         std::vector<iree_b200_command_buffer_t*>* cbs,
                                         const void* bias = nullptr) {
  iree_b200_device_t* device = ops->device;
  const CudaSymbols* cu = device->cu;
  const uint64_t es = matmul_elem_size(element_type);
  const uint64_t out_es = (flags & IREE_B200_MATMUL_FLAG_TRUNCATE_RESULT) ? 2 : 4;
  const bool transpose_rhs = (flags & IREE_B200_MATMUL_FLAG_TRANSPOSE_RHS) != 0;
  const bool accumulate = (flags & IREE_B200_MATMUL_FLAG_ACCUMULATE) != 0;
  // Uploading a fraction f of both operands enables only f^2 of the result: the download
  // stream idles early and runs behind late, whatever the panel order (the diagonal order
  // used here maximises the enabled area at every instant).
  const std::vector<int64_t> m_edges = panel_edges(M, ops->pipeline_block_m);
  const std::vector<int64_t> n_edges = panel_edges(N, ops->pipeline_block_n);
  const int64_t P = (int64_t)m_edges.size() - 1, Q = (int64_t)n_edges.size() - 1;
  const uint64_t off_a = 0;
  const uint64_t off_b = (bytes[0] + 1023) / 1024 * 1024;
  const uint64_t off_c = off_b + (bytes[1] + 1023) / 1024 * 1024;
  const uint64_t off_bias = off_c + (bytes[2] + 1023) / 1024 * 1024;
  const uint64_t bias_bytes = bias ? (uint64_t)N * 4 : 0;
  B200_RETURN_IF_ERROR(reserve_arena(ops, off_bias + bias_bytes + 1024));
  B200_RETURN_IF_ERROR(ensure_transfer_streams(ops));
  iree_b200_buffer_t* arena = ops->arena;
  const CUdeviceptr base = arena->device_ptr;
  if (bias) {
    // First on the upload stream: every panel's "ready" event is recorded behind it.
    B200_CU(cu, cuMemcpyHtoDAsync(base + off_bias, bias, bias_bytes, ops->upload_stream));
  }
  const uint8_t* host_a = static_cast<const uint8_t*>(lhs);
  const uint8_t* host_b = static_cast<const uint8_t*>(rhs);
  uint8_t* host_c = static_cast<uint8_t*>(out);
  size_t next_event = 0;
  std::vector<CUevent> a_ready((size_t)P), b_ready((size_t)Q);

  auto wait_on_dispatch_stream = [&](CUevent e) -> iree_b200_status_t {
    std::lock_guard<std::mutex> lock(device->mutex);
    B200_CU(cu, cuStreamWaitEvent(device->dispatch_stream, e, 0));
    return IREE_B200_STATUS_OK;
  };

  for (int64_t b = 0; b < batch; ++b) {
    const int64_t steps = std::max(P, Q);
    for (int64_t s = 0; s < steps; ++s) {
      if (s < P) {  // row panel s of A
        const int64_t rows = m_edges[s + 1] - m_edges[s];
        const uint64_t at = ((uint64_t)b * M + (uint64_t)m_edges[s]) * K * es;
        B200_RETURN_IF_ERROR(copy_2d(cu, true, base + off_a + at, const_cast<uint8_t*>(host_a) + at,
                                     K * es, K * es, rows, ops->upload_stream));
        B200_RETURN_IF_ERROR(event_at(ops, next_event++, &a_ready[s]));
        B200_CU(cu, cuEventRecord(a_ready[s], ops->upload_stream));
      }
      if (s < Q) {  // column panel s of B
        const int64_t cols = n_edges[s + 1] - n_edges[s];
        const uint64_t col0 = (uint64_t)n_edges[s];
        const uint64_t batch_at = (uint64_t)b * K * N * es;
        if (transpose_rhs) {
          const uint64_t at = batch_at + col0 * K * es;
          B200_RETURN_IF_ERROR(copy_2d(cu, true, base + off_b + at,
                                       const_cast<uint8_t*>(host_b) + at, K * es, K * es, cols,
                                       ops->upload_stream));
        } else {
          B200_RETURN_IF_ERROR(copy_2d(cu, true, base + off_b + batch_at + (uint64_t)K * col0 * es,
                                       const_cast<uint8_t*>(host_b) + batch_at + col0 * es,
                                       N * es, cols * es, K, ops->upload_stream));
        }
        B200_RETURN_IF_ERROR(event_at(ops, next_event++, &b_ready[s]));
        B200_CU(cu, cuEventRecord(b_ready[s], ops->upload_stream));
      }
      // Blocks enabled by this step: max(i, j) == s.
      for (int64_t i = 0; i <= std::min(s, P - 1); ++i) {
        for (int64_t j = 0; j <= std::min(s, Q - 1); ++j) {
          if (std::max(i, j) != s) continue;
          const int64_t rows = m_edges[i + 1] - m_edges[i], cols = n_edges[j + 1] - n_edges[j];
          const uint64_t row0 = (uint64_t)m_edges[i], col0 = (uint64_t)n_edges[j];
          const uint64_t a_at = ((uint64_t)b * M + row0) * K * es;
          const uint64_t b_at = (uint64_t)b * K * N * es + (uint64_t)K * col0 * es;
          // Staged result: row panel i is a run of dense [rows, cols_j] blocks.
          const uint64_t c_at = ((uint64_t)b * M * N + row0 * N + (uint64_t)rows * col0) * out_es;
          uint8_t* c_host = host_c + ((uint64_t)b * M * N + row0 * N + col0) * out_es;
          // The upload stream is in order (A_s, then B_s): the later panel's event covers both.
          CUevent ready = (j == s) ? b_ready[j] : a_ready[i];
          if (accumulate) {
            B200_RETURN_IF_ERROR(copy_2d(cu, true, base + off_c + c_at, c_host, N * out_es,
                                         cols * out_es, rows, ops->upload_stream));
            B200_RETURN_IF_ERROR(event_at(ops, next_event++, &ready));
            B200_CU(cu, cuEventRecord(ready, ops->upload_stream));
          }
          B200_RETURN_IF_ERROR(wait_on_dispatch_stream(ready));
          iree_b200_command_buffer_t* cb = nullptr;
          B200_RETURN_IF_ERROR(iree_b200_command_buffer_create(device, &cb));
          cbs->push_back(cb);
          B200_RETURN_IF_ERROR(iree_b200_command_buffer_begin(cb));
          const iree_b200_buffer_ref_t a_ref{arena, off_a + a_at, (uint64_t)rows * K * es};
          const iree_b200_buffer_ref_t b_ref{arena, off_b + b_at, (uint64_t)cols * K * es};
          const iree_b200_buffer_ref_t c_ref{arena, off_c + c_at, (uint64_t)rows * cols * out_es};
          if (bias) {  // this block's columns of the bias vector
            B200_RETURN_IF_ERROR(iree_b200_ops_record_matmul_bias(
                ops, cb, element_type, 1, rows, cols, K, flags, a_ref, b_ref,
                iree_b200_buffer_ref_t{arena, off_bias + col0 * 4, (uint64_t)cols * 4}, c_ref));
          } else {
            B200_RETURN_IF_ERROR(iree_b200_ops_record_matmul(ops, cb, element_type, 1, rows, cols, K,
                                                             flags, a_ref, b_ref, c_ref));
          }
          B200_RETURN_IF_ERROR(iree_b200_command_buffer_end(cb));
          B200_RETURN_IF_ERROR(
              iree_b200_device_queue_execute(device, 0, nullptr, nullptr, 0, nullptr, nullptr, cb));
          CUevent done = nullptr;
          B200_RETURN_IF_ERROR(event_at(ops, next_event++, &done));
          {
            std::lock_guard<std::mutex> lock(device->mutex);
            B200_CU(cu, cuEventRecord(done, device->dispatch_stream));
          }
          B200_CU(cu, cuStreamWaitEvent(ops->download_stream, done, 0));
          B200_RETURN_IF_ERROR(copy_2d(cu, false, base + off_c + c_at, c_host, N * out_es,
                                       cols * out_es, rows, ops->download_stream));
        }
      }
    }
  }
  B200_CU(cu, cuStreamSynchronize(ops->download_stream));
  return IREE_B200_STATUS_OK;
}

}  // namespace

static iree_b200_status_t matmul_host_impl(iree_b200_ops_t* ops, int32_t element_type,
                                           int64_t batch, int64_t M, int64_t N, int64_t K,
                                           uint32_t flags, const void* lhs, const void* rhs,
                                           const void* bias, void* out) {
  B200_REQUIRE_ARG(ops);
  B200_RETURN_IF_ERROR(require_device(ops));
  if (!matmul_type_ok(element_type) || batch < 0 || M < 0 || N < 0 || K < 0) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "bad matmul type or shape");
  }
  const uint64_t es = matmul_elem_size(element_type);
  const uint64_t out_es = (flags & IREE_B200_MATMUL_FLAG_TRUNCATE_RESULT) ? 2 : 4;
  const uint64_t bytes[3] = {(uint64_t)batch * M * K * es, (uint64_t)batch * K * N * es,
                             (uint64_t)batch * M * N * out_es};
  B200_REQUIRE_ARG((lhs || !bytes[0]) && (rhs || !bytes[1]) && (out || !bytes[2]));
  iree_b200_device_t* device = ops->device;
  std::lock_guard<std::mutex> host_lock(ops->host_mutex);
  // The pipeline needs every block to pass the TMA guards (16-byte pitches) and at least
  // two blocks to overlap; forced kernel families keep the single whole-problem dispatch.
  const bool tma_ok = K > 0 && (K * es) % 16 == 0 && (N * out_es) % 16 == 0 && (N * es) % 16 == 0;
  const int64_t blocks =
      batch * ceil_div(M, ops->pipeline_block_m) * ceil_div(N, ops->pipeline_block_n);
  // (With the threshold set to 0 — a test setting — a forced generic family is pipelined as
  // well: the mock-driver suite runs the SIMT kernels on the host and so checks the whole
  // panel / block data path without a GPU.)
  const bool family_forced =
      (flags & (IREE_B200_MATMUL_FLAG_FORCE_1CTA | IREE_B200_MATMUL_FLAG_FORCE_2CTA)) ||
      ((flags & IREE_B200_MATMUL_FLAG_FORCE_GENERIC) && ops->pipeline_min_bytes != 0);
  const bool pipelined = tma_ok && blocks >= 2 && !family_forced &&
                         bytes[0] + bytes[1] + bytes[2] >= ops->pipeline_min_bytes;
  std::vector<iree_b200_command_buffer_t*> cbs;
  iree_b200_status_t status =
      pipelined ? matmul_host_pipelined(ops, element_type, batch, M, N, K, flags, bytes, lhs, rhs,
                                        out, &cbs, bias)
                : matmul_host_serial(ops, element_type, batch, M, N, K, flags, bytes, lhs, rhs, out,
                                     &cbs, bias);
  std::string message = status != IREE_B200_STATUS_OK ? iree_b200_status_message() : "";
  if (status != IREE_B200_STATUS_OK) {
    // Drain everything that may still reference the staging arena or the host buffers.
    const CudaSymbols* cu = device->cu;
    if (ops->upload_stream) cu->cuStreamSynchronize(ops->upload_stream);
    iree_b200_device_wait_idle(device);
    if (ops->download_stream) cu->cuStreamSynchronize(ops->download_stream);
  }
  for (auto* cb : cbs) iree_b200_command_buffer_release(cb);
  if (status != IREE_B200_STATUS_OK) return make_status(status, "%s", message.c_str());
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t iree_b200_ops_matmul_host(iree_b200_ops_t* ops, int32_t element_type,
                                             int64_t batch, int64_t M, int64_t N, int64_t K,
                                             uint32_t flags, const void* lhs, const void* rhs,
                                             void* out) {
  if (flags & (IREE_B200_MATMUL_FLAG_BIAS | IREE_B200_MATMUL_FLAG_RELU)) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                       "FLAG_BIAS calls go through iree_b200_ops_matmul_bias_host");
  }
  return matmul_host_impl(ops, element_type, batch, M, N, K, flags, lhs, rhs, nullptr, out);
}

iree_b200_status_t iree_b200_ops_matmul_bias_host(iree_b200_ops_t* ops, int32_t element_type,
                                                  int64_t batch, int64_t M, int64_t N, int64_t K,
                                                  uint32_t flags, const void* lhs, const void* rhs,
                                                  const void* bias, void* out) {
  B200_REQUIRE_ARG(bias || !N);
  if (!bias) return IREE_B200_STATUS_OK;  // N == 0: empty result
  return matmul_host_impl(ops, element_type, batch, M, N, K,
                          flags | IREE_B200_MATMUL_FLAG_BIAS, lhs, rhs, bias, out);
}


//===----------------------------------------------------------------------===//
// The second builtin executable: kernels that have not had a GPU run yet (build.py
// EXPERIMENTAL_UNITS), loaded on first use so nothing validated depends on it.
// MXFP4 matmul (see kernels/mx_kernels.cu)
//===----------------------------------------------------------------------===//

static iree_b200_status_t ensure_pending_executable(iree_b200_ops_t* ops) {
  std::lock_guard<std::mutex> lock(ops->pending_mutex);  // not host_mutex: the host call holds that
  if (ops->pending_executable) return IREE_B200_STATUS_OK;
  // Everything is looked up into locals first; the executable and its ordinals are published
  // together only when all of them resolved, so a failed load can never leave a half
  // initialised table behind a non-null |pending_executable| (later calls would dispatch
  // ordinal -1).
  iree_b200_executable_t* executable = nullptr;
  B200_RETURN_IF_ERROR(iree_b200_executable_cache_prepare_executable(
      ops->cache, "PTXE", iree_b200_builtin_pending_cda1_start,
      (size_t)(iree_b200_builtin_pending_cda1_end - iree_b200_builtin_pending_cda1_start),
      &executable));
  static const char* const kNames[20] = {
      "iree_b200_matmul_generic_mxfp4_f32", "iree_b200_matmul_mxfp4_dequant_bf16",
      "iree_b200_mmt4d_generic_i8_i32",     "iree_b200_mmt4d_generic_f16_f32",
      "iree_b200_mmt4d_generic_bf16_f32",   "iree_b200_mmt4d_relayout_b8",
      "iree_b200_mmt4d_relayout_b16",       "iree_b200_mmt4d_relayout_b32",
      "iree_b200_matmul_mxfp4_f32",         "iree_b200_matmul_mxfp4_f32_1cta",
      "iree_b200_matmul_mxfp4_packed_f32",  "iree_b200_matmul_mxfp4_packed_f32_1cta",
      "iree_b200_mxfp4_pack_scales",        "iree_b200_mmt4d_generic_f32_f32",
      // tiles read directly: [slot I8 / F16 / BF16][2-CTA, 1-CTA]
      "iree_b200_mmt4d_i8_i32",             "iree_b200_mmt4d_i8_i32_1cta",
      "iree_b200_mmt4d_f16_f32",            "iree_b200_mmt4d_f16_f32_1cta",
      "iree_b200_mmt4d_bf16_f32",           "iree_b200_mmt4d_bf16_f32_1cta"};
  int32_t ordinals[20];
  for (int i = 0; i < 20; ++i) {
    ordinals[i] = lookup(executable, kNames[i]);
    if (ordinals[i] < 0) {
      iree_b200_executable_release(executable);
      return make_status(IREE_B200_STATUS_NOT_FOUND, "the second executable has no %s export",
                         kNames[i]);
    }
  }
  ops->mxfp4_generic_export = ordinals[0];
  ops->mxfp4_dequant_export = ordinals[1];
  ops->mxfp4_block_scaled_export = ordinals[8];
  ops->mxfp4_block_scaled_1cta_export = ordinals[9];
  ops->mxfp4_packed_export = ordinals[10];
  ops->mxfp4_packed_1cta_export = ordinals[11];
  ops->mxfp4_pack_scales_export = ordinals[12];
  ops->mmt4d_generic_exports[3] = ordinals[13];
  for (int i = 0; i < 6; ++i) ops->mmt4d_direct_exports[i / 2][i % 2] = ordinals[14 + i];
  for (int i = 0; i < 3; ++i) {
    ops->mmt4d_generic_exports[i] = ordinals[2 + i];
    ops->mmt4d_relayout_exports[i] = ordinals[5 + i];
  }
  ops->pending_executable = executable;
  return IREE_B200_STATUS_OK;
}

// The tensor-core route (expand both operands to bf16, then the bf16 tcgen05 NT GEMM) pays
// for shapes that fill at least one 128x128 tile and pass the GEMM's TMA guards; smaller or
// ragged shapes stay on the single SIMT dispatch.
static bool mxfp4_tensor_route(int64_t M, int64_t N, int64_t K) {
  return M >= 128 && N >= 128 && K >= 128 && (N * 4) % 16 == 0;
}

// Shapes the block-scaled tcgen05 kernel takes (K in whole 256-element stages, TMA-legal
// operand and result pitches).
static bool mxfp4_block_scaled_shape(int64_t M, int64_t N, int64_t K) {
  return M > 0 && N > 0 && K >= 256 && K % 256 == 0 && N % 4 == 0;
}

static uint64_t mxfp4_scale_atoms_bytes(int64_t rows, int64_t K) {
  return (uint64_t)ceil_div(rows, 128) * (uint64_t)(K / 128) * 512;
}

uint64_t iree_b200_ops_matmul_mxfp4_scale_workspace_size(int64_t M, int64_t N, int64_t K) {
  if (M < 0 || N < 0 || K < 0 || !mxfp4_block_scaled_shape(M, N, K)) return 0;
  return mxfp4_scale_atoms_bytes(M, K) + mxfp4_scale_atoms_bytes(N, K);  // multiples of 512
}

uint64_t iree_b200_ops_matmul_mxfp4_workspace_size(int64_t M, int64_t N, int64_t K) {
  if (M < 0 || N < 0 || K < 0 || !mxfp4_tensor_route(M, N, K)) return 0;
  return ((uint64_t)M * K * 2 + 255) / 256 * 256 + (uint64_t)N * K * 2;
}

static iree_b200_status_t record_matmul_mxfp4_impl(
    iree_b200_ops_t* ops, iree_b200_command_buffer_t* cb, int64_t M, int64_t N, int64_t K,
    uint32_t flags, iree_b200_buffer_ref_t lhs, iree_b200_buffer_ref_t lhs_scales,
    iree_b200_buffer_ref_t rhs, iree_b200_buffer_ref_t rhs_scales, iree_b200_buffer_ref_t out,
    const iree_b200_buffer_ref_t* workspace) {
  B200_REQUIRE_ARG(ops && cb);
  B200_RETURN_IF_ERROR(require_device(ops));
  if (flags & ~(uint32_t)(IREE_B200_MATMUL_FLAG_ACCUMULATE | IREE_B200_MATMUL_FLAG_FORCE_GENERIC |
                          IREE_B200_MATMUL_FLAG_MX_EXPAND | IREE_B200_MATMUL_FLAG_FORCE_1CTA |
                          IREE_B200_MATMUL_FLAG_FORCE_2CTA)) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                       "the MXFP4 matmul takes FLAG_ACCUMULATE, FLAG_FORCE_GENERIC and "
                       "FLAG_MX_EXPAND only (operands are always [rows, K])");
  }
  if (M < 0 || N < 0 || K < 0 || M > INT32_MAX || N > INT32_MAX || K > INT32_MAX ||
      K % IREE_B200_MX_BLOCK_SIZE) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                       "MXFP4 matmul shape %lldx%lldx%lld: K must be a multiple of the block size %d",
                       (long long)M, (long long)N, (long long)K, IREE_B200_MX_BLOCK_SIZE);
  }
  if (M == 0 || N == 0) return IREE_B200_STATUS_OK;  // empty result
  if (ceil_div(M, 64) > 65535) {
    return make_status(IREE_B200_STATUS_OUT_OF_RANGE, "MXFP4 matmul grid out of range");
  }
  const iree_b200_buffer_ref_t refs[5] = {lhs, lhs_scales, rhs, rhs_scales, out};
  const uint64_t need[5] = {(uint64_t)M * K / 2, (uint64_t)M * (K / IREE_B200_MX_BLOCK_SIZE),
                            (uint64_t)N * K / 2, (uint64_t)N * (K / IREE_B200_MX_BLOCK_SIZE),
                            (uint64_t)M * N * 4};
  for (int i = 0; i < 5; ++i) {
    // K == 0: the operand bindings are empty and may be NULL; the result is (out +) 0.
    if (!refs[i].buffer) {
      if (need[i]) return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "MXFP4 binding %d is NULL", i);
      continue;
    }
    const uint64_t total = refs[i].buffer->byte_length;
    if (refs[i].offset > total) {
      return make_status(IREE_B200_STATUS_OUT_OF_RANGE, "MXFP4 binding %d offset out of range", i);
    }
    const uint64_t length = refs[i].length == UINT64_MAX ? total - refs[i].offset : refs[i].length;
    if (length < need[i] || length > total - refs[i].offset) {
      return make_status(IREE_B200_STATUS_OUT_OF_RANGE,
                         "MXFP4 binding %d holds %llu bytes but the shape needs %llu", i,
                         (unsigned long long)length, (unsigned long long)need[i]);
    }
  }
  if ((out.buffer->device_ptr + out.offset) % 4) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "MXFP4 result must be 4-byte aligned");
  }
  B200_RETURN_IF_ERROR(ensure_pending_executable(ops));
  // The block-scaled tensor-core kernel (tcgen05.mma.kind::mxf4.block_scale): K in whole
  // 256-element stages, TMA-legal operand / result pitches and bases, 4-byte readable scale
  // rows. With a workspace the scales are first repacked into the instruction's atom order
  // (two small dispatches) and the kernel moves them by TMA (5.6 PFLOP/s at 8192^3); without,
  // one dispatch gathers them itself at about half that rate.
  auto aligned = [](const iree_b200_buffer_ref_t& r, uint64_t a) {
    return r.buffer && (r.buffer->device_ptr + r.offset) % a == 0;
  };
  if (!(flags & (IREE_B200_MATMUL_FLAG_FORCE_GENERIC | IREE_B200_MATMUL_FLAG_MX_EXPAND)) &&
      mxfp4_block_scaled_shape(M, N, K) && aligned(lhs, 16) && aligned(rhs, 16) && aligned(out, 16) &&
      aligned(lhs_scales, 4) && aligned(rhs_scales, 4)) {
    // Same family rule as the other tcgen05 GEMMs: SM pairs once there are enough 256 x 256
    // tiles to occupy them (FORCE_1CTA / FORCE_2CTA override).
    const int64_t sms = std::max<int64_t>(2, ops->sm_count - ops->reserved_sms);
    const int64_t tiles_2cta = ceil_div(M, 256) * ceil_div(N, 256);
    const int64_t tiles_1cta = ceil_div(M, 128) * ceil_div(N, 256);
    bool pair = M > 128 && tiles_2cta * 2 >= sms;
    if (flags & IREE_B200_MATMUL_FLAG_FORCE_1CTA) pair = false;
    if (flags & IREE_B200_MATMUL_FLAG_FORCE_2CTA) pair = true;
    const uint32_t grid[3] = {
        (uint32_t)(pair ? std::min<int64_t>(tiles_2cta * 2, sms & ~1) : std::min<int64_t>(tiles_1cta, sms)),
        1, 1};
    uint32_t kernel_flags = flags & IREE_B200_MATMUL_FLAG_ACCUMULATE;
    if (const char* diag = getenv("IREE_B200_MX_DIAG")) kernel_flags |= (uint32_t)strtoul(diag, nullptr, 0);
    const uint64_t sf_need = iree_b200_ops_matmul_mxfp4_scale_workspace_size(M, N, K);
    bool packed = false;
    if (workspace && workspace->buffer) {
      const uint64_t total = workspace->buffer->byte_length;
      if (workspace->offset > total) {
        return make_status(IREE_B200_STATUS_OUT_OF_RANGE, "MXFP4 workspace offset out of range");
      }
      const uint64_t length =
          workspace->length == UINT64_MAX ? total - workspace->offset : workspace->length;
      packed = length >= sf_need && length <= total - workspace->offset &&
               (workspace->buffer->device_ptr + workspace->offset) % 256 == 0;
    }
    if (packed) {
      const uint64_t a_atoms = mxfp4_scale_atoms_bytes(M, K), b_atoms = mxfp4_scale_atoms_bytes(N, K);
      const iree_b200_buffer_ref_t sfa{workspace->buffer, workspace->offset, a_atoms};
      const iree_b200_buffer_ref_t sfb{workspace->buffer, workspace->offset + a_atoms, b_atoms};
      const struct {
        iree_b200_buffer_ref_t scales, atoms;
        int64_t rows;
        uint64_t bytes;
      } sides[2] = {{lhs_scales, sfa, M, a_atoms}, {rhs_scales, sfb, N, b_atoms}};
      for (const auto& side : sides) {
        const uint32_t pgrid[3] = {(uint32_t)ceil_div((int64_t)(side.bytes / 4), 256), 1, 1};
        const uint32_t pconst[2] = {(uint32_t)side.rows, (uint32_t)K};
        const iree_b200_buffer_ref_t pbind[2] = {side.scales, side.atoms};
        B200_RETURN_IF_ERROR(iree_b200_command_buffer_dispatch(
            cb, ops->pending_executable, ops->mxfp4_pack_scales_export, pgrid, pconst, 2, pbind, 2, 0));
        ops->dispatch_count.fetch_add(1);
      }
      B200_RETURN_IF_ERROR(iree_b200_command_buffer_execution_barrier(cb));
      const uint32_t constants[8] = {(uint32_t)M,       (uint32_t)N,
                                     (uint32_t)K,       kernel_flags,
                                     (uint32_t)(K / 2), (uint32_t)(K / 128),
                                     (uint32_t)ceil_div(M, 128), (uint32_t)ceil_div(N, 128)};
      const iree_b200_buffer_ref_t packed_refs[5] = {lhs, sfa, rhs, sfb, out};
      B200_RETURN_IF_ERROR(iree_b200_command_buffer_dispatch(
          cb, ops->pending_executable, pair ? ops->mxfp4_packed_export : ops->mxfp4_packed_1cta_export,
          grid, constants, 8, packed_refs, 5, 0));
      ops->dispatch_count.fetch_add(1);
      return IREE_B200_STATUS_OK;
    }
    const uint32_t constants[5] = {(uint32_t)M, (uint32_t)N, (uint32_t)K, kernel_flags,
                                   (uint32_t)(K / 2)};
    B200_RETURN_IF_ERROR(iree_b200_command_buffer_dispatch(
        cb, ops->pending_executable,
        pair ? ops->mxfp4_block_scaled_export : ops->mxfp4_block_scaled_1cta_export, grid, constants,
        5, refs, 5, 0));
    ops->dispatch_count.fetch_add(1);
    return IREE_B200_STATUS_OK;
  }
  const uint64_t ws_need = iree_b200_ops_matmul_mxfp4_workspace_size(M, N, K);
  if (workspace && workspace->buffer && ws_need) {
    const uint64_t total = workspace->buffer->byte_length;
    if (workspace->offset > total) {
      return make_status(IREE_B200_STATUS_OUT_OF_RANGE, "MXFP4 workspace offset out of range");
    }
    const uint64_t length =
        workspace->length == UINT64_MAX ? total - workspace->offset : workspace->length;
    if (length < ws_need || length > total - workspace->offset ||
        (workspace->buffer->device_ptr + workspace->offset) % 256) {
      return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                         "MXFP4 workspace must be 256-byte aligned and hold %llu bytes "
                         "(iree_b200_ops_matmul_mxfp4_workspace_size)",
                         (unsigned long long)ws_need);
    }
    // Expand lhs and rhs to bf16 (exact, kernels/mx_kernels.cu), then the bf16 tcgen05 GEMM.
    const uint64_t a_bytes = (uint64_t)M * K * 2, b_bytes = (uint64_t)N * K * 2;
    const iree_b200_buffer_ref_t a_ws{workspace->buffer, workspace->offset, a_bytes};
    const iree_b200_buffer_ref_t b_ws{workspace->buffer,
                                      workspace->offset + (a_bytes + 255) / 256 * 256, b_bytes};
    const struct {
      iree_b200_buffer_ref_t packed, scales, expanded;
      int64_t rows;
    } sides[2] = {{lhs, lhs_scales, a_ws, M}, {rhs, rhs_scales, b_ws, N}};
    for (const auto& side : sides) {
      const int64_t blocks = ceil_div(side.rows * (K / 8), 256);
      if (blocks > INT32_MAX) {
        return make_status(IREE_B200_STATUS_OUT_OF_RANGE, "MXFP4 expansion grid out of range");
      }
      const uint32_t grid[3] = {(uint32_t)blocks, 1, 1};
      const uint32_t constants[2] = {(uint32_t)side.rows, (uint32_t)K};
      const iree_b200_buffer_ref_t bindings[3] = {side.packed, side.scales, side.expanded};
      B200_RETURN_IF_ERROR(iree_b200_command_buffer_dispatch(
          cb, ops->pending_executable, ops->mxfp4_dequant_export, grid, constants, 2, bindings, 3, 0));
      ops->dispatch_count.fetch_add(1);
    }
    B200_RETURN_IF_ERROR(iree_b200_command_buffer_execution_barrier(cb));
    return iree_b200_ops_record_matmul(
        ops, cb, IREE_B200_ELEMENT_TYPE_BF16, 1, M, N, K,
        IREE_B200_MATMUL_FLAG_TRANSPOSE_RHS | (flags & IREE_B200_MATMUL_FLAG_ACCUMULATE), a_ws, b_ws,
        out);
  }
  const uint32_t grid[3] = {(uint32_t)ceil_div(N, 64), (uint32_t)ceil_div(M, 64), 1};
  const uint32_t constants[5] = {(uint32_t)M, (uint32_t)N, (uint32_t)K, 1u,
                                 flags & IREE_B200_MATMUL_FLAG_ACCUMULATE};
  B200_RETURN_IF_ERROR(iree_b200_command_buffer_dispatch(cb, ops->pending_executable,
                                                         ops->mxfp4_generic_export, grid, constants,
                                                         5, refs, 5, 0));
  ops->dispatch_count.fetch_add(1);
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t iree_b200_ops_record_matmul_mxfp4(
    iree_b200_ops_t* ops, iree_b200_command_buffer_t* cb, int64_t M, int64_t N, int64_t K,
    uint32_t flags, iree_b200_buffer_ref_t lhs, iree_b200_buffer_ref_t lhs_scales,
    iree_b200_buffer_ref_t rhs, iree_b200_buffer_ref_t rhs_scales, iree_b200_buffer_ref_t out) {
  return record_matmul_mxfp4_impl(ops, cb, M, N, K, flags, lhs, lhs_scales, rhs, rhs_scales, out,
                                  nullptr);
}

iree_b200_status_t iree_b200_ops_record_matmul_mxfp4_tensor(
    iree_b200_ops_t* ops, iree_b200_command_buffer_t* cb, int64_t M, int64_t N, int64_t K,
    uint32_t flags, iree_b200_buffer_ref_t lhs, iree_b200_buffer_ref_t lhs_scales,
    iree_b200_buffer_ref_t rhs, iree_b200_buffer_ref_t rhs_scales, iree_b200_buffer_ref_t workspace,
    iree_b200_buffer_ref_t out) {
  return record_matmul_mxfp4_impl(ops, cb, M, N, K, flags, lhs, lhs_scales, rhs, rhs_scales, out,
                                  &workspace);
}

iree_b200_status_t iree_b200_ops_matmul_mxfp4_host(iree_b200_ops_t* ops, int64_t M, int64_t N,
                                                   int64_t K, uint32_t flags, const void* lhs,
                                                   const void* lhs_scales, const void* rhs,
                                                   const void* rhs_scales, void* out) {
  B200_REQUIRE_ARG(ops);
  B200_RETURN_IF_ERROR(require_device(ops));
  if (M < 0 || N < 0 || K < 0 || K % IREE_B200_MX_BLOCK_SIZE) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "bad MXFP4 matmul shape");
  }
  const uint64_t bytes[5] = {(uint64_t)M * K / 2, (uint64_t)M * (K / IREE_B200_MX_BLOCK_SIZE),
                             (uint64_t)N * K / 2, (uint64_t)N * (K / IREE_B200_MX_BLOCK_SIZE),
                             (uint64_t)M * N * 4};
  const void* host_in[4] = {lhs, lhs_scales, rhs, rhs_scales};
  for (int i = 0; i < 4; ++i) B200_REQUIRE_ARG(host_in[i] || !bytes[i]);
  B200_REQUIRE_ARG(out || !bytes[4]);
  if (!bytes[4]) return IREE_B200_STATUS_OK;
  B200_RETURN_IF_ERROR(ensure_pending_executable(ops));
  iree_b200_device_t* device = ops->device;
  std::lock_guard<std::mutex> host_lock(ops->host_mutex);
  uint64_t offsets[5], total = 0;
  for (int i = 0; i < 5; ++i) {
    offsets[i] = total;
    total += (bytes[i] + 1023) / 1024 * 1024;
  }
  // The expansion workspace is only needed when the bf16-expansion route can be taken.
  // Workspace of the route the call will take: packed scale atoms for the block-scaled
  // kernel, the bf16 expansion of both operands for the expansion route, none for SIMT.
  const bool block_scaled = !(flags & IREE_B200_MATMUL_FLAG_MX_EXPAND) && mxfp4_block_scaled_shape(M, N, K);
  const uint64_t ws_offset = total,
                 ws_bytes = (flags & IREE_B200_MATMUL_FLAG_FORCE_GENERIC) ? 0
                            : block_scaled ? iree_b200_ops_matmul_mxfp4_scale_workspace_size(M, N, K)
                                           : iree_b200_ops_matmul_mxfp4_workspace_size(M, N, K);
  total += ws_bytes;
  iree_b200_command_buffer_t* cb = nullptr;
  auto run = [&]() -> iree_b200_status_t {
    B200_RETURN_IF_ERROR(reserve_arena(ops, total + 1024));
    auto ref = [&](int i) {
      return bytes[i] ? iree_b200_buffer_ref_t{ops->arena, offsets[i], bytes[i]}
                      : iree_b200_buffer_ref_t{nullptr, 0, 0};
    };
    const iree_b200_buffer_ref_t workspace =
        ws_bytes ? iree_b200_buffer_ref_t{ops->arena, ws_offset, ws_bytes}
                 : iree_b200_buffer_ref_t{nullptr, 0, 0};
    for (int i = 0; i < 4; ++i) {
      if (!bytes[i]) continue;
      B200_RETURN_IF_ERROR(
          iree_b200_device_transfer_h2d(device, host_in[i], ops->arena, offsets[i], bytes[i]));
    }
    if (flags & IREE_B200_MATMUL_FLAG_ACCUMULATE) {
      B200_RETURN_IF_ERROR(iree_b200_device_transfer_h2d(device, out, ops->arena, offsets[4], bytes[4]));
    }
    B200_RETURN_IF_ERROR(iree_b200_command_buffer_create(device, &cb));
    B200_RETURN_IF_ERROR(iree_b200_command_buffer_begin(cb));
    B200_RETURN_IF_ERROR(iree_b200_ops_record_matmul_mxfp4_tensor(
        ops, cb, M, N, K, flags, ref(0), ref(1), ref(2), ref(3), workspace, ref(4)));
    B200_RETURN_IF_ERROR(iree_b200_command_buffer_end(cb));
    B200_RETURN_IF_ERROR(
        iree_b200_device_queue_execute(device, 0, nullptr, nullptr, 0, nullptr, nullptr, cb));
    B200_RETURN_IF_ERROR(iree_b200_device_transfer_d2h(device, ops->arena, offsets[4], out, bytes[4]));
    return iree_b200_device_wait_idle(device);
  };
  iree_b200_status_t status = run();
  std::string message = status != IREE_B200_STATUS_OK ? iree_b200_status_message() : "";
  if (status != IREE_B200_STATUS_OK) iree_b200_device_wait_idle(device);
  iree_b200_command_buffer_release(cb);
  if (status != IREE_B200_STATUS_OK) return make_status(status, "%s", message.c_str());
  return IREE_B200_STATUS_OK;
}

//===----------------------------------------------------------------------===//
// linalg.mmt4d on data-tiled operands (pending its first GPU run; kernels/mmt4d_kernels.cu)
//===----------------------------------------------------------------------===//

namespace {

struct Mmt4dShape {
  int slot;             // index into mmt4d_generic_exports
  uint64_t bytes[3];    // lhs, rhs, out
};

iree_b200_status_t check_mmt4d(int32_t element_type, int64_t M1, int64_t N1, int64_t K1, int32_t M0,
                               int32_t N0, int32_t K0, uint32_t flags, Mmt4dShape* shape) {
  switch (element_type) {
    case IREE_B200_ELEMENT_TYPE_I8: shape->slot = 0; break;
    case IREE_B200_ELEMENT_TYPE_F16: shape->slot = 1; break;
    case IREE_B200_ELEMENT_TYPE_BF16: shape->slot = 2; break;
    case IREE_B200_ELEMENT_TYPE_F32: shape->slot = 3; break;  // SIMT only (no exact f32 tensor-core form)
    default:
      return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                         "mmt4d takes i8, f16, bf16 or f32 operands (got element type %d)", element_type);
  }
  if (flags & ~(uint32_t)(IREE_B200_MATMUL_FLAG_ACCUMULATE | IREE_B200_MATMUL_FLAG_FORCE_GENERIC |
                          IREE_B200_MATMUL_FLAG_MMT4D_RELAYOUT)) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                       "mmt4d takes FLAG_ACCUMULATE and FLAG_FORCE_GENERIC only");
  }
  if (M0 < 1 || N0 < 1 || K0 < 1 || M0 > 32767 || N0 > 32767 || K0 > 32767) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                       "mmt4d tile sizes %dx%dx%d must be within [1, 32767]", M0, N0, K0);
  }
  // The logical extents index with 32-bit ints in the kernel; rows are grid.y blocks of 64.
  if (M1 < 0 || N1 < 0 || K1 < 0 || M1 > INT32_MAX / M0 || N1 > INT32_MAX / N0 ||
      K1 > INT32_MAX / K0 || ceil_div(M1 * M0, 64) > 65535) {
    return make_status(IREE_B200_STATUS_OUT_OF_RANGE,
                       "mmt4d shape [%lld,%lld,%lld] x tiles %dx%dx%d out of range", (long long)M1,
                       (long long)N1, (long long)K1, M0, N0, K0);
  }
  const uint64_t es = element_type == IREE_B200_ELEMENT_TYPE_F32 ? 4 : matmul_elem_size(element_type);
  shape->bytes[0] = (uint64_t)M1 * K1 * M0 * K0 * es;
  shape->bytes[1] = (uint64_t)N1 * K1 * N0 * K0 * es;
  shape->bytes[2] = (uint64_t)M1 * N1 * M0 * N0 * 4;
  return IREE_B200_STATUS_OK;
}

}  // namespace

// The tensor-core route: unpack lhs / rhs (and the accumulator) to row-major in a workspace,
// run the tcgen05 NT GEMM, pack the result back into tiles. Worth it once the logical shape
// fills a 128x128 tile and passes the GEMM's TMA guards.
static uint64_t round256(uint64_t v) { return (v + 255) / 256 * 256; }

uint64_t iree_b200_ops_mmt4d_workspace_size(int32_t element_type, int64_t M1, int64_t N1, int64_t K1,
                                            int32_t M0, int32_t N0, int32_t K0) {
  Mmt4dShape shape;
  if (check_mmt4d(element_type, M1, N1, K1, M0, N0, K0, 0, &shape) != IREE_B200_STATUS_OK) return 0;
  if (element_type == IREE_B200_ELEMENT_TYPE_F32) return 0;  // SIMT only
  const int64_t M = M1 * M0, N = N1 * N0, K = K1 * K0;
  const int64_t es = (int64_t)matmul_elem_size(element_type);
  if (M < 128 || N < 128 || K < 128 || (K * es) % 16 || (N * 4) % 16) return 0;
  return round256(shape.bytes[0]) + round256(shape.bytes[1]) + shape.bytes[2];
}

static iree_b200_status_t record_mmt4d_relayout(iree_b200_ops_t* ops, iree_b200_command_buffer_t* cb,
                                                uint64_t elem_size, int64_t R1, int64_t C1,
                                                int32_t R0, int32_t C0, bool to_tiled,
                                                iree_b200_buffer_ref_t src,
                                                iree_b200_buffer_ref_t dst) {
  // One work item per (tile row, panel of tiles that fits the kernel's 16 KiB staging buffer),
  // strided over a 1-D grid of at most 8 blocks per SM (mmt4d_kernels.cu).
  const int64_t tile = (int64_t)R0 * C0, panel_elems = 16384 / (int64_t)elem_size;
  const int64_t items = tile > panel_elems ? ceil_div(R1 * R0 * C1 * C0, 256)
                                           : R1 * ceil_div(C1, std::max<int64_t>(1, panel_elems / tile));
  const uint32_t grid[3] = {(uint32_t)std::max<int64_t>(1, std::min<int64_t>(items, 8 * (int64_t)ops->sm_count)), 1, 1};
  const uint32_t constants[5] = {(uint32_t)R1, (uint32_t)C1, (uint32_t)R0, (uint32_t)C0,
                                 to_tiled ? 1u : 0u};
  const iree_b200_buffer_ref_t bindings[2] = {src, dst};
  const int slot = elem_size == 1 ? 0 : elem_size == 2 ? 1 : 2;
  B200_RETURN_IF_ERROR(iree_b200_command_buffer_dispatch(
      cb, ops->pending_executable, ops->mmt4d_relayout_exports[slot], grid, constants, 5, bindings, 2,
      0));
  ops->dispatch_count.fetch_add(1);
  return IREE_B200_STATUS_OK;
}

// Shapes the tile-direct tcgen05 exports take: K0 rows of 16 bytes, M0 / N0 multiples of 8
// dividing the 128-row boxes (8 tile rows are one UMMA core matrix), a problem that fills a
// tile.
static bool mmt4d_direct_shape(int32_t element_type, int64_t M1, int64_t N1, int64_t K1, int32_t M0,
                               int32_t N0, int32_t K0, uint32_t flags) {
  if (element_type == IREE_B200_ELEMENT_TYPE_F32 ||
      (flags & (IREE_B200_MATMUL_FLAG_FORCE_GENERIC | IREE_B200_MATMUL_FLAG_MMT4D_RELAYOUT))) {
    return false;
  }
  const int64_t es = (int64_t)matmul_elem_size(element_type);
  return K0 * es == 16 && 128 % M0 == 0 && 128 % N0 == 0 && M0 % 8 == 0 && N0 % 8 == 0 &&
         M1 * M0 >= 128 && N1 * N0 >= 128 && K1 * K0 >= 128;
}

static iree_b200_status_t record_mmt4d_impl(iree_b200_ops_t* ops, iree_b200_command_buffer_t* cb,
                                            int32_t element_type, int64_t M1, int64_t N1,
                                            int64_t K1, int32_t M0, int32_t N0, int32_t K0,
                                            uint32_t flags, iree_b200_buffer_ref_t lhs,
                                            iree_b200_buffer_ref_t rhs, iree_b200_buffer_ref_t out,
                                            const iree_b200_buffer_ref_t* workspace) {
  B200_REQUIRE_ARG(ops && cb);
  B200_RETURN_IF_ERROR(require_device(ops));
  Mmt4dShape shape;
  B200_RETURN_IF_ERROR(check_mmt4d(element_type, M1, N1, K1, M0, N0, K0, flags, &shape));
  if (M1 == 0 || N1 == 0) return IREE_B200_STATUS_OK;  // empty result
  // K1 == 0 with ACCUMULATE leaves the result as it is (iree_uk_mmt4d_early, mmt4d.c:106-114).
  if (K1 == 0 && (flags & IREE_B200_MATMUL_FLAG_ACCUMULATE)) return IREE_B200_STATUS_OK;
  const iree_b200_buffer_ref_t refs[3] = {lhs, rhs, out};
  const uint64_t in_align = element_type == IREE_B200_ELEMENT_TYPE_F32 ? 4 : matmul_elem_size(element_type);
  const uint64_t align[3] = {in_align, in_align, 4};
  for (int i = 0; i < 3; ++i) {
    if (!refs[i].buffer) {  // K1 == 0: the operand bindings are empty and may be NULL
      if (shape.bytes[i]) return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "mmt4d binding %d is NULL", i);
      continue;
    }
    const uint64_t total = refs[i].buffer->byte_length;
    if (refs[i].offset > total) {
      return make_status(IREE_B200_STATUS_OUT_OF_RANGE, "mmt4d binding %d offset out of range", i);
    }
    const uint64_t length = refs[i].length == UINT64_MAX ? total - refs[i].offset : refs[i].length;
    if (length < shape.bytes[i] || length > total - refs[i].offset) {
      return make_status(IREE_B200_STATUS_OUT_OF_RANGE,
                         "mmt4d binding %d holds %llu bytes but the shape needs %llu", i,
                         (unsigned long long)length, (unsigned long long)shape.bytes[i]);
    }
    if ((refs[i].buffer->device_ptr + refs[i].offset) % align[i]) {
      return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                         "mmt4d binding %d is not aligned to its element size", i);
    }
  }
  B200_RETURN_IF_ERROR(ensure_pending_executable(ops));
  // Tiles read and written directly by the tcgen05 GEMM (one dispatch, no workspace): K0 rows
  // of 16 bytes, M0 / N0 multiples of 8 dividing the 128-row boxes (8 tile rows are one UMMA
  // core matrix), 16-byte aligned bindings, a problem that fills a tile.
  {
    const int64_t M = M1 * M0, N = N1 * N0;
    bool direct = mmt4d_direct_shape(element_type, M1, N1, K1, M0, N0, K0, flags);
    for (int i = 0; i < 3 && direct; ++i) {
      direct = (refs[i].buffer->device_ptr + refs[i].offset) % 16 == 0;
    }
    if (direct) {
      const int64_t sms = std::max<int64_t>(2, ops->sm_count - ops->reserved_sms);
      const int64_t tiles_2cta = ceil_div(M, 256) * ceil_div(N, 256);
      const int64_t tiles_1cta = ceil_div(M, 128) * ceil_div(N, 256);
      const bool pair = M > 128 && tiles_2cta * 2 >= sms;
      const uint32_t grid[3] = {
          (uint32_t)(pair ? std::min<int64_t>(tiles_2cta * 2, sms & ~1) : std::min<int64_t>(tiles_1cta, sms)),
          1, 1};
      const uint32_t c_cols = (uint32_t)std::min<int32_t>(N0, 32);  // result box row, in elements
      const uint32_t c_swizzle = c_cols == 32 ? 3u : c_cols == 16 ? 2u : 1u;  // 128 / 64 / 32 B
      const uint32_t constants[14] = {(uint32_t)M1, (uint32_t)N1, (uint32_t)K1, (uint32_t)M0,
                                      (uint32_t)N0, (uint32_t)K0,
                                      (flags & IREE_B200_MATMUL_FLAG_ACCUMULATE) |
                                          (getenv("IREE_B200_MMT4D_DIAG")
                                               ? (uint32_t)strtoul(getenv("IREE_B200_MMT4D_DIAG"), nullptr, 0)
                                               : 0u),
                                      (uint32_t)(M0 / 8), (uint32_t)(128 / M0), (uint32_t)(N0 / 8),
                                      (uint32_t)((pair ? 128 : 256) / N0), c_cols, 32 / c_cols,
                                      c_swizzle};
      B200_RETURN_IF_ERROR(iree_b200_command_buffer_dispatch(
          cb, ops->pending_executable, ops->mmt4d_direct_exports[shape.slot][pair ? 0 : 1], grid,
          constants, 14, refs, 3, 0));
      ops->dispatch_count.fetch_add(1);
      return IREE_B200_STATUS_OK;
    }
  }
  const uint64_t ws_need =
      iree_b200_ops_mmt4d_workspace_size(element_type, M1, N1, K1, M0, N0, K0);
  if (workspace && workspace->buffer && ws_need) {
    const uint64_t total = workspace->buffer->byte_length;
    if (workspace->offset > total) {
      return make_status(IREE_B200_STATUS_OUT_OF_RANGE, "mmt4d workspace offset out of range");
    }
    const uint64_t length =
        workspace->length == UINT64_MAX ? total - workspace->offset : workspace->length;
    if (length < ws_need || length > total - workspace->offset ||
        (workspace->buffer->device_ptr + workspace->offset) % 256) {
      return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                         "mmt4d workspace must be 256-byte aligned and hold %llu bytes "
                         "(iree_b200_ops_mmt4d_workspace_size)",
                         (unsigned long long)ws_need);
    }
    const uint64_t es = matmul_elem_size(element_type);
    const uint64_t a_at = workspace->offset, b_at = a_at + round256(shape.bytes[0]),
                   c_at = b_at + round256(shape.bytes[1]);
    const iree_b200_buffer_ref_t a_rm{workspace->buffer, a_at, shape.bytes[0]};
    const iree_b200_buffer_ref_t b_rm{workspace->buffer, b_at, shape.bytes[1]};
    const iree_b200_buffer_ref_t c_rm{workspace->buffer, c_at, shape.bytes[2]};
    const bool accumulate = (flags & IREE_B200_MATMUL_FLAG_ACCUMULATE) != 0;
    B200_RETURN_IF_ERROR(record_mmt4d_relayout(ops, cb, es, M1, K1, M0, K0, false, lhs, a_rm));
    B200_RETURN_IF_ERROR(record_mmt4d_relayout(ops, cb, es, N1, K1, N0, K0, false, rhs, b_rm));
    if (accumulate) {
      B200_RETURN_IF_ERROR(record_mmt4d_relayout(ops, cb, 4, M1, N1, M0, N0, false, out, c_rm));
    }
    B200_RETURN_IF_ERROR(iree_b200_command_buffer_execution_barrier(cb));
    B200_RETURN_IF_ERROR(iree_b200_ops_record_matmul(
        ops, cb, element_type, 1, M1 * M0, N1 * N0, K1 * K0,
        IREE_B200_MATMUL_FLAG_TRANSPOSE_RHS | (flags & IREE_B200_MATMUL_FLAG_ACCUMULATE), a_rm, b_rm,
        c_rm));
    B200_RETURN_IF_ERROR(iree_b200_command_buffer_execution_barrier(cb));
    return record_mmt4d_relayout(ops, cb, 4, M1, N1, M0, N0, true, c_rm, out);
  }
  const uint32_t grid[3] = {(uint32_t)ceil_div(N1 * N0, 64), (uint32_t)ceil_div(M1 * M0, 64), 1};
  const uint32_t constants[7] = {(uint32_t)M1, (uint32_t)N1, (uint32_t)K1, (uint32_t)M0,
                                 (uint32_t)N0, (uint32_t)K0,
                                 flags & IREE_B200_MATMUL_FLAG_ACCUMULATE};
  B200_RETURN_IF_ERROR(iree_b200_command_buffer_dispatch(
      cb, ops->pending_executable, ops->mmt4d_generic_exports[shape.slot], grid, constants, 7, refs, 3, 0));
  ops->dispatch_count.fetch_add(1);
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t iree_b200_ops_record_mmt4d(iree_b200_ops_t* ops, iree_b200_command_buffer_t* cb,
                                              int32_t element_type, int64_t M1, int64_t N1,
                                              int64_t K1, int32_t M0, int32_t N0, int32_t K0,
                                              uint32_t flags, iree_b200_buffer_ref_t lhs,
                                              iree_b200_buffer_ref_t rhs, iree_b200_buffer_ref_t out) {
  return record_mmt4d_impl(ops, cb, element_type, M1, N1, K1, M0, N0, K0, flags, lhs, rhs, out,
                           nullptr);
}

iree_b200_status_t iree_b200_ops_record_mmt4d_tensor(
    iree_b200_ops_t* ops, iree_b200_command_buffer_t* cb, int32_t element_type, int64_t M1,
    int64_t N1, int64_t K1, int32_t M0, int32_t N0, int32_t K0, uint32_t flags,
    iree_b200_buffer_ref_t lhs, iree_b200_buffer_ref_t rhs, iree_b200_buffer_ref_t workspace,
    iree_b200_buffer_ref_t out) {
  return record_mmt4d_impl(ops, cb, element_type, M1, N1, K1, M0, N0, K0, flags, lhs, rhs, out,
                           &workspace);
}

iree_b200_status_t iree_b200_ops_mmt4d_host(iree_b200_ops_t* ops, int32_t element_type, int64_t M1,
                                            int64_t N1, int64_t K1, int32_t M0, int32_t N0,
                                            int32_t K0, uint32_t flags, const void* lhs,
                                            const void* rhs, void* out) {
  B200_REQUIRE_ARG(ops);
  B200_RETURN_IF_ERROR(require_device(ops));
  Mmt4dShape shape;
  B200_RETURN_IF_ERROR(check_mmt4d(element_type, M1, N1, K1, M0, N0, K0, flags, &shape));
  const uint64_t* bytes = shape.bytes;
  B200_REQUIRE_ARG((lhs || !bytes[0]) && (rhs || !bytes[1]) && (out || !bytes[2]));
  const bool accumulate = (flags & IREE_B200_MATMUL_FLAG_ACCUMULATE) != 0;
  if (!bytes[2] || (K1 == 0 && accumulate)) return IREE_B200_STATUS_OK;
  iree_b200_device_t* device = ops->device;
  std::lock_guard<std::mutex> host_lock(ops->host_mutex);
  uint64_t offsets[3], total = 0;
  for (int i = 0; i < 3; ++i) {
    offsets[i] = total;
    total += (bytes[i] + 1023) / 1024 * 1024;
  }
  // FORCE_GENERIC keeps the call on the single SIMT dispatch (k-ascending accumulation).
  const uint64_t ws_offset = total,
                 ws_bytes = ((flags & IREE_B200_MATMUL_FLAG_FORCE_GENERIC) ||
                             mmt4d_direct_shape(element_type, M1, N1, K1, M0, N0, K0, flags))
                                ? 0
                                : iree_b200_ops_mmt4d_workspace_size(element_type, M1, N1, K1, M0, N0, K0);
  total += ws_bytes;
  iree_b200_command_buffer_t* cb = nullptr;
  auto run = [&]() -> iree_b200_status_t {
    B200_RETURN_IF_ERROR(reserve_arena(ops, total + 1024));
    auto ref = [&](int i) {
      return bytes[i] ? iree_b200_buffer_ref_t{ops->arena, offsets[i], bytes[i]}
                      : iree_b200_buffer_ref_t{nullptr, 0, 0};
    };
    const void* host_in[2] = {lhs, rhs};
    for (int i = 0; i < 2; ++i) {
      if (!bytes[i]) continue;
      B200_RETURN_IF_ERROR(
          iree_b200_device_transfer_h2d(device, host_in[i], ops->arena, offsets[i], bytes[i]));
    }
    if (accumulate) {
      B200_RETURN_IF_ERROR(iree_b200_device_transfer_h2d(device, out, ops->arena, offsets[2], bytes[2]));
    }
    B200_RETURN_IF_ERROR(iree_b200_command_buffer_create(device, &cb));
    B200_RETURN_IF_ERROR(iree_b200_command_buffer_begin(cb));
    const iree_b200_buffer_ref_t workspace =
        ws_bytes ? iree_b200_buffer_ref_t{ops->arena, ws_offset, ws_bytes}
                 : iree_b200_buffer_ref_t{nullptr, 0, 0};
    B200_RETURN_IF_ERROR(iree_b200_ops_record_mmt4d_tensor(
        ops, cb, element_type, M1, N1, K1, M0, N0, K0, flags, ref(0), ref(1), workspace, ref(2)));
    B200_RETURN_IF_ERROR(iree_b200_command_buffer_end(cb));
    B200_RETURN_IF_ERROR(
        iree_b200_device_queue_execute(device, 0, nullptr, nullptr, 0, nullptr, nullptr, cb));
    B200_RETURN_IF_ERROR(iree_b200_device_transfer_d2h(device, ops->arena, offsets[2], out, bytes[2]));
    return iree_b200_device_wait_idle(device);
  };
  iree_b200_status_t status = run();
  std::string message = status != IREE_B200_STATUS_OK ? iree_b200_status_message() : "";
  if (status != IREE_B200_STATUS_OK) iree_b200_device_wait_idle(device);
  iree_b200_command_buffer_release(cb);
  if (status != IREE_B200_STATUS_OK) return make_status(status, "%s", message.c_str());
  return IREE_B200_STATUS_OK;
}

//===----------------------------------------------------------------------===//
// iree_linalg_ext.attention
//===----------------------------------------------------------------------===//

static uint32_t float_bits(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  return u;
}

iree_b200_status_t iree_b200_ops_select_attention(const iree_b200_ops_t* ops, int32_t element_type,
                                                  int64_t B, int64_t M, int64_t K1, int64_t K2,
                                                  int64_t N, float scale, int32_t mask_kind,
                                                  uint32_t flags, uint32_t alignment,
                                                  iree_b200_attention_selection_t* out_selection) {
  B200_REQUIRE_ARG(ops && out_selection);
  memset(out_selection, 0, sizeof(*out_selection));
  if (element_type != IREE_B200_ELEMENT_TYPE_F16 && element_type != IREE_B200_ELEMENT_TYPE_BF16 &&
      element_type != IREE_B200_ELEMENT_TYPE_F32) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                       "attention element type %d unsupported (f16 / bf16 / f32)", element_type);
  }
  if (mask_kind < 0 || mask_kind > IREE_B200_ATTENTION_MASK_CAUSAL) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "bad attention mask kind %d", mask_kind);
  }
  if (mask_kind == IREE_B200_ATTENTION_MASK_CAUSAL) {
    if (K2 < M) {
      return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                         "causal attention needs K2 >= M (row m sees keys <= m + K2 - M); got "
                         "M=%lld K2=%lld", (long long)M, (long long)K2);
    }
    // A single query row sees every key: the mask is all-true.
    if (M == 1) mask_kind = IREE_B200_ATTENTION_MASK_NONE;
  }
  const int64_t dims[5] = {B, M, K1, K2, N};
  for (int64_t d : dims) {
    if (d < 0 || d > INT32_MAX) {
      return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "attention dimension %lld out of range",
                         (long long)d);
    }
  }
  if (B == 0 || M == 0 || N == 0) return IREE_B200_STATUS_OK;  // empty result
  if (K1 == 0 || K2 == 0) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                       "attention with an empty reduction (K1=%lld, K2=%lld) is undefined",
                       (long long)K1, (long long)K2);
  }
  // The tcgen05 prefill kernel takes no mask, the causal mask (computed in the kernel) or an
  // i1 mask operand whose rows can be read 16 bytes at a time; the decode kernels take none.
  const bool tc_operands = K1 == 128 && N == 128 && alignment >= 16 &&
                           element_type != IREE_B200_ELEMENT_TYPE_F32;
  const bool prefill_mask_ok =
      mask_kind == IREE_B200_ATTENTION_MASK_NONE || mask_kind == IREE_B200_ATTENTION_MASK_CAUSAL ||
      (mask_kind == IREE_B200_ATTENTION_MASK_BOOL && K2 % 16 == 0);
  const bool tc_shape = tc_operands && (M > 16 ? prefill_mask_ok
                                               : mask_kind == IREE_B200_ATTENTION_MASK_NONE);
  enum { kPrefill, kDecode, kGeneric } family;
  if (flags & IREE_B200_ATTENTION_FLAG_FORCE_GENERIC) {
    family = kGeneric;
  } else if (flags & (IREE_B200_ATTENTION_FLAG_FORCE_PREFILL | IREE_B200_ATTENTION_FLAG_FORCE_DECODE |
                      IREE_B200_ATTENTION_FLAG_FORCE_DECODE_SPLIT |
                      IREE_B200_ATTENTION_FLAG_FORCE_DECODE_CLUSTER)) {
    family = (flags & IREE_B200_ATTENTION_FLAG_FORCE_PREFILL) ? kPrefill : kDecode;
    if (!tc_operands || (family == kPrefill ? !prefill_mask_ok
                                            : mask_kind != IREE_B200_ATTENTION_MASK_NONE)) {
      return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                         "shape/mask/alignment does not satisfy the guards of the forced kernel");
    }
  } else if (tc_shape) {
    family = M > 16 ? kPrefill : kDecode;
  } else {
    family = kGeneric;
  }
  const uint32_t scale_bits = float_bits(scale);
  iree_b200_attention_dispatch_t* d0 = &out_selection->dispatches[0];
  iree_b200_attention_dispatch_t* d1 = &out_selection->dispatches[1];
  switch (family) {
    case kPrefill: {
      if (B > 65535) return make_status(IREE_B200_STATUS_OUT_OF_RANGE, "attention batch too large");
      int variant = kDefaultPrefillVariant;
      if (flags & IREE_B200_ATTENTION_FLAG_EXP2_MUFU_ONLY) variant = kAttnPrefill;
      if (flags & IREE_B200_ATTENTION_FLAG_EXP2_POLY_25) variant = kAttnPrefillPoly25;
      if (flags & IREE_B200_ATTENTION_FLAG_EXP2_POLY_50) variant = kAttnPrefillPoly50;
      // Tail split: all units (two Q tiles of one batch slice) cost the same, so the last
      // wave of the grid is only partly full unless the unit count divides by the SM count.
      // Cut just that wave's units into |split| CTAs over K2 (+ one small merge dispatch)
      // when that makes the last wave at most half as long.
      const int64_t tiles_m = ceil_div(M, 256);
      const int64_t units = tiles_m * B;
      const int64_t sms = ops->sm_count;
      const int64_t kv_blocks = ceil_div(K2, 128);
      const int64_t full_units = units / sms * sms;
      const int64_t tail_units = units - full_units;
      int64_t split = tail_units ? std::min<int64_t>(4, sms / tail_units) : 1;
      split = std::min(split, kv_blocks / 4);  // every part keeps >= 4 K2 blocks
      const bool device_maps = (flags & IREE_B200_ATTENTION_FLAG_DEVICE_MAPS) != 0;
      if (device_maps) {
        // The export an unmodified IREE CUDA driver can launch: one dispatch, no mask tensor.
        if (mask_kind == IREE_B200_ATTENTION_MASK_BOOL || variant != kAttnPrefill) {
          return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                             "FLAG_DEVICE_MAPS: unmasked or causal prefill, default exp2 only");
        }
        variant = kAttnPrefillDeviceMaps;
      }
      const bool use_split = !(flags & IREE_B200_ATTENTION_FLAG_NO_TAIL_SPLIT) && split >= 2 &&
                             variant == kAttnPrefill && mask_kind == IREE_B200_ATTENTION_MASK_NONE &&
                             full_units + tail_units * split <= 0x7fffffff;
      if (use_split) {
        d0->export_ordinal = ops->attention_exports[element_type][kAttnPrefillSplit];
        d0->workgroup_count[0] = (uint32_t)(full_units + tail_units * split);
        d0->workgroup_count[1] = 1;
        d0->workgroup_count[2] = 1;
        d0->constant_count = 6;
        d0->constants[0] = (uint32_t)B;
        d0->constants[1] = (uint32_t)M;
        d0->constants[2] = (uint32_t)K2;
        d0->constants[3] = scale_bits;
        d0->constants[4] = (uint32_t)full_units;
        d0->constants[5] = (uint32_t)split;
        d1->export_ordinal = ops->attention_exports[element_type][kAttnPrefillMerge];
        d1->workgroup_count[0] = 32;  // 256 rows of a unit, 8 per CTA
        d1->workgroup_count[1] = (uint32_t)tail_units;
        d1->workgroup_count[2] = 1;
        d1->constant_count = 5;
        d1->constants[0] = (uint32_t)B;
        d1->constants[1] = (uint32_t)M;
        d1->constants[2] = (uint32_t)full_units;
        d1->constants[3] = (uint32_t)split;
        d1->constants[4] = (uint32_t)tail_units;
        out_selection->dispatch_count = 2;
        out_selection->workspace_bytes = (uint64_t)tail_units * split * 256 * (128 + 2) * 4;
        break;
      }
      if (mask_kind == IREE_B200_ATTENTION_MASK_BOOL) variant = kAttnPrefillMask;
      d0->export_ordinal = ops->attention_exports[element_type][variant];
      d0->workgroup_count[0] = (uint32_t)ceil_div(M, 256);
      d0->workgroup_count[1] = (uint32_t)B;
      d0->workgroup_count[2] = 1;
      d0->constant_count = 5;
      d0->constants[0] = (uint32_t)B;
      d0->constants[1] = (uint32_t)M;
      d0->constants[2] = (uint32_t)K2;
      d0->constants[3] = scale_bits;
      d0->constants[4] = 0;
      if (mask_kind == IREE_B200_ATTENTION_MASK_CAUSAL) {
        // Units differ in work (row m walks K2 blocks up to its limit only): a 1-D grid that
        // the kernel reads heaviest-first, so the hardware's in-order block dispatch is
        // longest-processing-time-first scheduling.
        if (units > 0x7fffffff) return make_status(IREE_B200_STATUS_OUT_OF_RANGE, "attention grid too large");
        d0->workgroup_count[0] = (uint32_t)units;
        d0->workgroup_count[1] = 1;
        d0->constants[4] = 1;  // kAttnFlagCausal
      }
      out_selection->dispatch_count = 1;
      break;
    }
    case kDecode: {
      if (B > 65535 || M > 65535) {
        return make_status(IREE_B200_STATUS_OUT_OF_RANGE, "attention decode grid out of range");
      }
      using Cluster = b200::AttnDecodeClusterConfig;
      // Single dispatch when 8 CTAs per (batch, m) row already fill the SMs and every CTA
      // gets a few ring stages of rows; otherwise split-KV sized to the SM count + merge.
      bool cluster = B * M * Cluster::kSplits >= ops->sm_count &&
                     K2 >= (int64_t)Cluster::kSplits * Cluster::kMinRowsPerSplit;
      if (flags & IREE_B200_ATTENTION_FLAG_FORCE_DECODE_SPLIT) cluster = false;
      if (flags & IREE_B200_ATTENTION_FLAG_FORCE_DECODE_CLUSTER) cluster = true;
      if (cluster) {
        d0->export_ordinal = ops->attention_exports[element_type][kAttnDecodeCluster];
        d0->workgroup_count[0] = Cluster::kSplits;
        d0->workgroup_count[1] = (uint32_t)M;
        d0->workgroup_count[2] = (uint32_t)B;
        d0->constant_count = 5;
        d0->constants[0] = (uint32_t)B;
        d0->constants[1] = (uint32_t)M;
        d0->constants[2] = (uint32_t)K2;
        d0->constants[3] = scale_bits;
        d0->constants[4] = 0;
        out_selection->dispatch_count = 1;
        break;
      }
      const int64_t chunk = b200::AttnDecodeConfig::kChunkRows;
      const int64_t want = std::max<int64_t>(1, (int64_t)ops->sm_count * 8 / (B * M));
      const int64_t splits = std::max<int64_t>(1, std::min(ceil_div(K2, chunk), want));
      d0->export_ordinal = ops->attention_exports[element_type][kAttnDecodeSplit];
      d0->workgroup_count[0] = (uint32_t)splits;
      d0->workgroup_count[1] = (uint32_t)M;
      d0->workgroup_count[2] = (uint32_t)B;
      d0->constant_count = 5;
      d0->constants[0] = (uint32_t)B;
      d0->constants[1] = (uint32_t)M;
      d0->constants[2] = (uint32_t)K2;
      d0->constants[3] = scale_bits;
      d0->constants[4] = (uint32_t)splits;
      d1->export_ordinal = ops->attention_exports[element_type][kAttnDecodeMerge];
      d1->workgroup_count[0] = (uint32_t)M;
      d1->workgroup_count[1] = (uint32_t)B;
      d1->workgroup_count[2] = 1;
      d1->constant_count = 3;
      d1->constants[0] = (uint32_t)B;
      d1->constants[1] = (uint32_t)M;
      d1->constants[2] = (uint32_t)splits;
      out_selection->dispatch_count = 2;
      out_selection->workspace_bytes =
          (uint64_t)B * M * splits * b200::AttnDecodeConfig::kRecordFloats * 4;
      break;
    }
    default: {
      if (K1 > b200::AttnGenericConfig::kMaxK1 || N > b200::AttnGenericConfig::kMaxN) {
        return make_status(IREE_B200_STATUS_UNIMPLEMENTED,
                           "attention K1=%lld N=%lld exceeds the generic kernel's limits; such "
                           "ops stay with the compiler's own codegen",
                           (long long)K1, (long long)N);
      }
      if (B > 65535) return make_status(IREE_B200_STATUS_OUT_OF_RANGE, "attention batch too large");
      d0->export_ordinal = ops->attention_exports[element_type][kAttnGeneric];
      d0->workgroup_count[0] = (uint32_t)M;
      d0->workgroup_count[1] = (uint32_t)B;
      d0->workgroup_count[2] = 1;
      d0->constant_count = 7;
      d0->constants[0] = (uint32_t)B;
      d0->constants[1] = (uint32_t)M;
      d0->constants[2] = (uint32_t)K1;
      d0->constants[3] = (uint32_t)K2;
      d0->constants[4] = (uint32_t)N;
      d0->constants[5] = scale_bits;
      d0->constants[6] = (uint32_t)mask_kind;
      out_selection->dispatch_count = 1;
      break;
    }
  }
  for (int i = 0; i < out_selection->dispatch_count; ++i) {
    if (out_selection->dispatches[i].export_ordinal < 0) {
      return make_status(IREE_B200_STATUS_NOT_FOUND,
                         "the executable lacks an attention export for element type %d",
                         element_type);
    }
  }
  return IREE_B200_STATUS_OK;
}

static iree_b200_status_t check_binding(const char* what, iree_b200_buffer_ref_t ref, uint64_t need,
                                        uint32_t* alignment) {
  if (!ref.buffer) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "attention binding %s is null", what);
  }
  const uint64_t total = ref.buffer->byte_length;
  if (ref.offset > total) {
    return make_status(IREE_B200_STATUS_OUT_OF_RANGE, "attention binding %s offset out of range",
                       what);
  }
  const uint64_t length = ref.length == UINT64_MAX ? total - ref.offset : ref.length;
  if (length < need || length > total - ref.offset) {
    return make_status(IREE_B200_STATUS_OUT_OF_RANGE,
                       "attention binding %s holds %llu bytes but the shape needs %llu", what,
                       (unsigned long long)length, (unsigned long long)need);
  }
  *alignment = std::min(*alignment, pointer_alignment(ref.buffer->device_ptr + ref.offset));
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t iree_b200_ops_record_attention(
    iree_b200_ops_t* ops, iree_b200_command_buffer_t* cb, int32_t element_type, int64_t B,
    int64_t M, int64_t K1, int64_t K2, int64_t N, float scale, int32_t mask_kind, uint32_t flags,
    iree_b200_buffer_ref_t q, iree_b200_buffer_ref_t k, iree_b200_buffer_ref_t v,
    iree_b200_buffer_ref_t mask, iree_b200_buffer_ref_t workspace, iree_b200_buffer_ref_t out) {
  B200_REQUIRE_ARG(ops && cb);
  B200_RETURN_IF_ERROR(require_device(ops));
  if (B < 0 || M < 0 || K1 < 0 || K2 < 0 || N < 0) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "negative attention dimension");
  }
  uint32_t alignment = 256;
  const uint64_t es = element_type == IREE_B200_ELEMENT_TYPE_F32 ? 4 : 2;
  B200_RETURN_IF_ERROR(check_binding("Q", q, (uint64_t)B * M * K1 * es, &alignment));
  B200_RETURN_IF_ERROR(check_binding("K", k, (uint64_t)B * K2 * K1 * es, &alignment));
  B200_RETURN_IF_ERROR(check_binding("V", v, (uint64_t)B * K2 * N * es, &alignment));
  B200_RETURN_IF_ERROR(check_binding("O", out, (uint64_t)B * M * N * es, &alignment));
  if (mask_kind == IREE_B200_ATTENTION_MASK_BOOL || mask_kind == IREE_B200_ATTENTION_MASK_F32) {
    // The i1 mask of the tcgen05 prefill kernel is read 16 bytes at a time: its alignment
    // counts like the operands'; the generic kernel reads it element by element.
    uint32_t mask_alignment = 256;
    B200_RETURN_IF_ERROR(check_binding(
        "mask", mask, (uint64_t)B * M * K2 * (mask_kind == IREE_B200_ATTENTION_MASK_F32 ? 4 : 1),
        &mask_alignment));
    if (mask_kind == IREE_B200_ATTENTION_MASK_BOOL) alignment = std::min(alignment, mask_alignment);
  }
  iree_b200_attention_selection_t sel;
  B200_RETURN_IF_ERROR(iree_b200_ops_select_attention(ops, element_type, B, M, K1, K2, N, scale,
                                                      mask_kind, flags, alignment, &sel));
  if (sel.dispatch_count == 0) return IREE_B200_STATUS_OK;
  if (sel.workspace_bytes) {
    uint32_t ignored = 256;
    B200_RETURN_IF_ERROR(check_binding("workspace", workspace, sel.workspace_bytes, &ignored));
  }
  if (sel.dispatch_count == 2) {
    // Decode split: (Q, K, V, workspace) then (workspace, O). Prefill tail split:
    // (Q, K, V, O, workspace) then (workspace, O).
    const bool prefill_split =
        sel.dispatches[0].export_ordinal == ops->attention_exports[element_type][kAttnPrefillSplit];
    const iree_b200_buffer_ref_t decode_refs[4] = {q, k, v, workspace};
    const iree_b200_buffer_ref_t prefill_refs[5] = {q, k, v, out, workspace};
    B200_RETURN_IF_ERROR(iree_b200_command_buffer_dispatch(
        cb, ops->executable, sel.dispatches[0].export_ordinal, sel.dispatches[0].workgroup_count,
        sel.dispatches[0].constants, sel.dispatches[0].constant_count,
        prefill_split ? prefill_refs : decode_refs, prefill_split ? 5 : 4, 0));
    B200_RETURN_IF_ERROR(iree_b200_command_buffer_execution_barrier(cb));
    const iree_b200_buffer_ref_t merge_refs[2] = {workspace, out};
    B200_RETURN_IF_ERROR(iree_b200_command_buffer_dispatch(
        cb, ops->executable, sel.dispatches[1].export_ordinal, sel.dispatches[1].workgroup_count,
        sel.dispatches[1].constants, sel.dispatches[1].constant_count, merge_refs, 2, 0));
    ops->dispatch_count.fetch_add(2);
    return IREE_B200_STATUS_OK;
  }
  if (element_type != IREE_B200_ELEMENT_TYPE_F32 &&
      sel.dispatches[0].export_ordinal == ops->attention_exports[element_type][kAttnPrefillMask]) {
    const iree_b200_buffer_ref_t refs[5] = {q, k, v, mask, out};
    B200_RETURN_IF_ERROR(iree_b200_command_buffer_dispatch(
        cb, ops->executable, sel.dispatches[0].export_ordinal, sel.dispatches[0].workgroup_count,
        sel.dispatches[0].constants, 5, refs, 5, 0));
  } else if (sel.dispatches[0].constant_count == 7) {
    // Generic export: the mask binding must be a valid pointer even when unused.
    const bool has_operand = mask_kind == IREE_B200_ATTENTION_MASK_BOOL ||
                             mask_kind == IREE_B200_ATTENTION_MASK_F32;
    const iree_b200_buffer_ref_t refs[5] = {q, k, v, has_operand ? mask : q, out};
    B200_RETURN_IF_ERROR(iree_b200_command_buffer_dispatch(
        cb, ops->executable, sel.dispatches[0].export_ordinal, sel.dispatches[0].workgroup_count,
        sel.dispatches[0].constants, 7, refs, 5, 0));
  } else {
    const iree_b200_buffer_ref_t refs[4] = {q, k, v, out};
    B200_RETURN_IF_ERROR(iree_b200_command_buffer_dispatch(
        cb, ops->executable, sel.dispatches[0].export_ordinal, sel.dispatches[0].workgroup_count,
        sel.dispatches[0].constants, 5, refs, 4, 0));
  }
  ops->dispatch_count.fetch_add(1);
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t iree_b200_ops_attention_host(iree_b200_ops_t* ops, int32_t element_type,
                                                int64_t B, int64_t M, int64_t K1, int64_t K2,
                                                int64_t N, float scale, int32_t mask_kind,
                                                uint32_t flags, const void* q, const void* k,
                                                const void* v, const void* mask, void* out) {
  B200_REQUIRE_ARG(ops);
  B200_RETURN_IF_ERROR(require_device(ops));
  if (B < 0 || M < 0 || K1 < 0 || K2 < 0 || N < 0 || mask_kind < 0 ||
      mask_kind > IREE_B200_ATTENTION_MASK_CAUSAL) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "bad attention shape or mask kind");
  }
  iree_b200_attention_selection_t sel;
  B200_RETURN_IF_ERROR(iree_b200_ops_select_attention(ops, element_type, B, M, K1, K2, N, scale,
                                                      mask_kind, flags, 256, &sel));
  if (sel.dispatch_count == 0) return IREE_B200_STATUS_OK;
  const uint64_t mask_bytes =
      (mask_kind == IREE_B200_ATTENTION_MASK_BOOL || mask_kind == IREE_B200_ATTENTION_MASK_F32)
          ? (uint64_t)B * M * K2 * (mask_kind == IREE_B200_ATTENTION_MASK_F32 ? 4 : 1)
          : 0;  // none, or the causal mask the kernels compute themselves
  // 0=Q 1=K 2=V 3=mask 4=workspace 5=O
  const uint64_t es = element_type == IREE_B200_ELEMENT_TYPE_F32 ? 4 : 2;
  const uint64_t bytes[6] = {(uint64_t)B * M * K1 * es, (uint64_t)B * K2 * K1 * es,
                             (uint64_t)B * K2 * N * es,  mask_bytes,
                             sel.workspace_bytes,        (uint64_t)B * M * N * es};
  const void* host_in[4] = {q, k, v, mask};
  for (int i = 0; i < 4; ++i) B200_REQUIRE_ARG(host_in[i] || !bytes[i]);
  B200_REQUIRE_ARG(out);
  iree_b200_device_t* device = ops->device;
  std::lock_guard<std::mutex> host_lock(ops->host_mutex);
  uint64_t offsets[6], total = 0;
  for (int i = 0; i < 6; ++i) {
    offsets[i] = total;
    total += (bytes[i] + 1023) / 1024 * 1024;
  }
  // Batch entries are independent, and every operand is [B, ...] dense: above the pipeline
  // threshold the call runs in slices of B on the three streams of the matmul pipeline
  // (slice c+1 goes up while slice c computes and slice c-1 comes down), so the call costs
  // about the upload alone. Decode-shaped calls (M <= 16) stay whole: their kernel choice
  // depends on B and there is nothing to overlap but a 28 us dispatch.
  const uint64_t in_bytes = bytes[0] + bytes[1] + bytes[2] + bytes[3];
  const uint64_t per_entry_in = in_bytes / (uint64_t)B;
  // About half the pipeline threshold per slice (8 MiB by default: long enough for PCIe to
  // run at its full rate).
  const uint64_t slice_bytes = std::max<uint64_t>(1, ops->pipeline_min_bytes / 2);
  int64_t slice =
      (int64_t)std::max<uint64_t>(1, ceil_div(slice_bytes, std::max<uint64_t>(1, per_entry_in)));
  slice = std::max<int64_t>(slice, ceil_div(B, 64));
  const bool pipelined = M > 16 && B >= 2 * slice && in_bytes + bytes[5] >= ops->pipeline_min_bytes;
  std::vector<iree_b200_command_buffer_t*> cbs;
  if (pipelined) {
    // Slices share one workspace (they run in order on the dispatch stream); size it for
    // the largest request (whole slices and the remainder select differently).
    uint64_t workspace = 0;
    for (int64_t entries : {slice, B % slice}) {
      if (!entries) continue;
      iree_b200_attention_selection_t s;
      B200_RETURN_IF_ERROR(iree_b200_ops_select_attention(ops, element_type, entries, M, K1, K2, N,
                                                          scale, mask_kind, flags, 256, &s));
      workspace = std::max<uint64_t>(workspace, s.workspace_bytes);
    }
    total += (workspace + 1023) / 1024 * 1024;  // appended after O; offsets[4] is unused here
  }
  auto run_pipelined = [&]() -> iree_b200_status_t {
    const CudaSymbols* cu = device->cu;
    B200_RETURN_IF_ERROR(reserve_arena(ops, total + 1024));
    B200_RETURN_IF_ERROR(ensure_transfer_streams(ops));
    const CUdeviceptr base = ops->arena->device_ptr;
    const uint64_t off_workspace = offsets[5] + (bytes[5] + 1023) / 1024 * 1024;
    const uint64_t workspace_bytes = total - off_workspace;
    size_t next_event = 0;
    for (int64_t b0 = 0; b0 < B; b0 += slice) {
      const int64_t entries = std::min(slice, B - b0);
      uint64_t at[6], len[6];
      for (int i = 0; i < 6; ++i) {
        const uint64_t per_entry = bytes[i] / (uint64_t)B;
        at[i] = offsets[i] + (uint64_t)b0 * per_entry;
        len[i] = (uint64_t)entries * per_entry;
      }
      for (int i = 0; i < 4; ++i) {
        if (!len[i]) continue;
        B200_CU(cu, cuMemcpyHtoDAsync(base + at[i],
                                      static_cast<const uint8_t*>(host_in[i]) + (at[i] - offsets[i]),
                                      len[i], ops->upload_stream));
      }
      CUevent landed = nullptr, done = nullptr;
      B200_RETURN_IF_ERROR(event_at(ops, next_event++, &landed));
      B200_CU(cu, cuEventRecord(landed, ops->upload_stream));
      {
        std::lock_guard<std::mutex> lock(device->mutex);
        B200_CU(cu, cuStreamWaitEvent(device->dispatch_stream, landed, 0));
      }
      auto part = [&](int i) {
        return len[i] ? iree_b200_buffer_ref_t{ops->arena, at[i], len[i]}
                      : iree_b200_buffer_ref_t{nullptr, 0, 0};
      };
      iree_b200_command_buffer_t* cb = nullptr;
      B200_RETURN_IF_ERROR(iree_b200_command_buffer_create(device, &cb));
      cbs.push_back(cb);
      B200_RETURN_IF_ERROR(iree_b200_command_buffer_begin(cb));
      B200_RETURN_IF_ERROR(iree_b200_ops_record_attention(
          ops, cb, element_type, entries, M, K1, K2, N, scale, mask_kind, flags, part(0), part(1),
          part(2), part(3),
          workspace_bytes ? iree_b200_buffer_ref_t{ops->arena, off_workspace, workspace_bytes}
                          : iree_b200_buffer_ref_t{nullptr, 0, 0},
          part(5)));
      B200_RETURN_IF_ERROR(iree_b200_command_buffer_end(cb));
      B200_RETURN_IF_ERROR(
          iree_b200_device_queue_execute(device, 0, nullptr, nullptr, 0, nullptr, nullptr, cb));
      B200_RETURN_IF_ERROR(event_at(ops, next_event++, &done));
      {
        std::lock_guard<std::mutex> lock(device->mutex);
        B200_CU(cu, cuEventRecord(done, device->dispatch_stream));
      }
      B200_CU(cu, cuStreamWaitEvent(ops->download_stream, done, 0));
      B200_CU(cu, cuMemcpyDtoHAsync(static_cast<uint8_t*>(out) + (at[5] - offsets[5]), base + at[5],
                                    len[5], ops->download_stream));
    }
    B200_CU(cu, cuStreamSynchronize(ops->download_stream));
    return IREE_B200_STATUS_OK;
  };
  iree_b200_command_buffer_t* cb = nullptr;
  auto ref = [&](int i) {
    return bytes[i] ? iree_b200_buffer_ref_t{ops->arena, offsets[i], bytes[i]}
                    : iree_b200_buffer_ref_t{nullptr, 0, 0};
  };
  auto run = [&]() -> iree_b200_status_t {
    B200_RETURN_IF_ERROR(reserve_arena(ops, total + 1024));
    for (int i = 0; i < 4; ++i) {
      if (!bytes[i]) continue;
      B200_RETURN_IF_ERROR(
          iree_b200_device_transfer_h2d(device, host_in[i], ops->arena, offsets[i], bytes[i]));
    }
    B200_RETURN_IF_ERROR(iree_b200_command_buffer_create(device, &cb));
    B200_RETURN_IF_ERROR(iree_b200_command_buffer_begin(cb));
    B200_RETURN_IF_ERROR(iree_b200_ops_record_attention(ops, cb, element_type, B, M, K1, K2, N,
                                                        scale, mask_kind, flags, ref(0), ref(1),
                                                        ref(2), ref(3), ref(4), ref(5)));
    B200_RETURN_IF_ERROR(iree_b200_command_buffer_end(cb));
    B200_RETURN_IF_ERROR(
        iree_b200_device_queue_execute(device, 0, nullptr, nullptr, 0, nullptr, nullptr, cb));
    B200_RETURN_IF_ERROR(iree_b200_device_transfer_d2h(device, ops->arena, offsets[5], out, bytes[5]));
    return iree_b200_device_wait_idle(device);
  };
  iree_b200_status_t status = pipelined ? run_pipelined() : run();
  std::string message = status != IREE_B200_STATUS_OK ? iree_b200_status_message() : "";
  if (status != IREE_B200_STATUS_OK) {
    // Drain everything that may still reference the staging arena or the host buffers.
    const CudaSymbols* cu = device->cu;
    if (ops->upload_stream) cu->cuStreamSynchronize(ops->upload_stream);
    iree_b200_device_wait_idle(device);
    if (ops->download_stream) cu->cuStreamSynchronize(ops->download_stream);
  }
  iree_b200_command_buffer_release(cb);
  for (auto* c : cbs) iree_b200_command_buffer_release(c);
  if (status != IREE_B200_STATUS_OK) return make_status(status, "%s", message.c_str());
  return IREE_B200_STATUS_OK;
}

}  // extern "C"
