// ops.cc — export selection ("matcher guards + count region") and dispatch recording for
// the dense-contraction ops. See include/iree_b200_ops.h.

#include <algorithm>

#include "../hal/internal.h"
#include "iree_b200_ops.h"

using namespace b200hal;

extern "C" {
// Provided by builtin_blob.S (.incbin of the packed CDA1 executable).
extern const uint8_t iree_b200_builtin_cda1_start[];
extern const uint8_t iree_b200_builtin_cda1_end[];
}

namespace {

enum MatmulFamily { kFamily2Cta = 0, kFamily1Cta = 1, kFamilyGeneric = 2, kFamilyCount = 3 };

const char* const kTypeSuffix[3] = {"i8_i32", "f16_f32", "bf16_f32"};

}  // namespace

struct iree_b200_ops_t {
  iree_b200_device_t* device = nullptr;
  iree_b200_executable_cache_t* cache = nullptr;
  iree_b200_executable_t* executable = nullptr;
  std::atomic<uint64_t> dispatch_count{0};
  // [element_type][transpose_rhs][family] → export ordinal (-1 if absent)
  int32_t matmul_exports[3][2][kFamilyCount];
  int32_t attention_exports[3][8];
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
  for (int t = 0; t < 3; ++t) {
    for (int nt = 0; nt < 2; ++nt) {
      const std::string base = std::string("iree_b200_matmul_") + (nt ? "nt_" : "") + kTypeSuffix[t];
      ops->matmul_exports[t][nt][kFamily2Cta] = lookup(ops->executable, base);
      ops->matmul_exports[t][nt][kFamily1Cta] = lookup(ops->executable, base + "_1cta");
      ops->matmul_exports[t][nt][kFamilyGeneric] =
          lookup(ops->executable, std::string("iree_b200_matmul_generic_") + kTypeSuffix[t]);
    }
  }
  *out_ops = ops;
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t iree_b200_ops_create(iree_b200_device_t* device, iree_b200_ops_t** out_ops) {
  size_t size = 0;
  const void* data = iree_b200_ops_builtin_executable_data(&size);
  return iree_b200_ops_create_from_data(device, data, size, out_ops);
}

void iree_b200_ops_release(iree_b200_ops_t* ops) {
  if (!ops) return;
  iree_b200_executable_release(ops->executable);
  iree_b200_executable_cache_release(ops->cache);
  iree_b200_device_release(ops->device);
  delete ops;
}

iree_b200_executable_t* iree_b200_ops_executable(iree_b200_ops_t* ops) {
  return ops ? ops->executable : nullptr;
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
  if (element_type < 0 || element_type > 2) {
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
  const int64_t es = element_type == IREE_B200_ELEMENT_TYPE_I8 ? 1 : 2;
  out_selection->constants[0] = (uint32_t)M;
  out_selection->constants[1] = (uint32_t)N;
  out_selection->constants[2] = (uint32_t)K;
  out_selection->constants[3] = (uint32_t)batch;
  out_selection->constants[4] = flags & IREE_B200_MATMUL_FLAG_ACCUMULATE;
  out_selection->export_ordinal = -1;
  if (batch == 0 || M == 0 || N == 0) return IREE_B200_STATUS_OK;  // empty result: no dispatch

  // Matcher guards of the TMA-fed kernels: 16-byte aligned bases and row pitches.
  const bool tma_ok = alignment >= 16 && K > 0 && (K * es) % 16 == 0 && (N * 4) % 16 == 0 &&
                      (transpose_rhs || (N * es) % 16 == 0);
  const int sm_count = ops->device->sm_count;
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
  const int32_t ordinal = ops->matmul_exports[element_type][transpose_rhs ? 1 : 0][family];
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
      out_selection->constants[4] |= transpose_rhs ? 2u : 0u;
      break;
  }
  return IREE_B200_STATUS_OK;
}

static uint32_t pointer_alignment(uint64_t p) {
  uint32_t a = 256;
  while (a > 1 && (p % a) != 0) a >>= 1;
  return a;
}

iree_b200_status_t iree_b200_ops_record_matmul(iree_b200_ops_t* ops, iree_b200_command_buffer_t* cb,
                                               int32_t element_type, int64_t batch, int64_t M,
                                               int64_t N, int64_t K, uint32_t flags,
                                               iree_b200_buffer_ref_t lhs,
                                               iree_b200_buffer_ref_t rhs,
                                               iree_b200_buffer_ref_t out) {
  B200_REQUIRE_ARG(ops && cb);
  B200_REQUIRE_ARG(lhs.buffer && rhs.buffer && out.buffer);
  const uint64_t es = element_type == IREE_B200_ELEMENT_TYPE_I8 ? 1 : 2;
  // Binding extents must cover the operands (the HAL validation layer's job,
  // hal/command_buffer_validation.c).
  const uint64_t need[3] = {(uint64_t)batch * M * K * es, (uint64_t)batch * K * N * es,
                            (uint64_t)batch * M * N * 4};
  const iree_b200_buffer_ref_t refs[3] = {lhs, rhs, out};
  uint32_t alignment = 256;
  for (int i = 0; i < 3; ++i) {
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
    alignment = std::min(alignment,
                         pointer_alignment(refs[i].buffer->device_ptr + refs[i].offset));
  }
  iree_b200_matmul_selection_t sel;
  B200_RETURN_IF_ERROR(
      iree_b200_ops_select_matmul(ops, element_type, batch, M, N, K, flags, alignment, &sel));
  if (sel.export_ordinal < 0) return IREE_B200_STATUS_OK;  // empty result
  B200_RETURN_IF_ERROR(iree_b200_command_buffer_dispatch(cb, ops->executable, sel.export_ordinal,
                                                         sel.workgroup_count, sel.constants, 5, refs,
                                                         3, 0));
  ops->dispatch_count.fetch_add(1);
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t iree_b200_ops_matmul_host(iree_b200_ops_t* ops, int32_t element_type,
                                             int64_t batch, int64_t M, int64_t N, int64_t K,
                                             uint32_t flags, const void* lhs, const void* rhs,
                                             void* out) {
  B200_REQUIRE_ARG(ops);
  if (element_type < 0 || element_type > 2 || batch < 0 || M < 0 || N < 0 || K < 0) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "bad matmul type or shape");
  }
  const uint64_t es = element_type == IREE_B200_ELEMENT_TYPE_I8 ? 1 : 2;
  const uint64_t bytes[3] = {(uint64_t)batch * M * K * es, (uint64_t)batch * K * N * es,
                             (uint64_t)batch * M * N * 4};
  B200_REQUIRE_ARG((lhs || !bytes[0]) && (rhs || !bytes[1]) && (out || !bytes[2]));
  iree_b200_device_t* device = ops->device;
  iree_b200_buffer_t* buffers[3] = {nullptr, nullptr, nullptr};
  iree_b200_command_buffer_t* cb = nullptr;
  iree_b200_status_t status = IREE_B200_STATUS_OK;
  auto run = [&]() -> iree_b200_status_t {
    for (int i = 0; i < 3; ++i) {
      B200_RETURN_IF_ERROR(iree_b200_allocator_allocate_buffer(
          device, IREE_B200_MEMORY_TYPE_DEVICE_LOCAL, bytes[i], &buffers[i]));
    }
    B200_RETURN_IF_ERROR(iree_b200_device_transfer_h2d(device, lhs, buffers[0], 0, bytes[0]));
    B200_RETURN_IF_ERROR(iree_b200_device_transfer_h2d(device, rhs, buffers[1], 0, bytes[1]));
    if (flags & IREE_B200_MATMUL_FLAG_ACCUMULATE) {
      B200_RETURN_IF_ERROR(iree_b200_device_transfer_h2d(device, out, buffers[2], 0, bytes[2]));
    }
    B200_RETURN_IF_ERROR(iree_b200_command_buffer_create(device, &cb));
    B200_RETURN_IF_ERROR(iree_b200_command_buffer_begin(cb));
    if (K == 0 && !(flags & IREE_B200_MATMUL_FLAG_ACCUMULATE) && bytes[2]) {
      // Empty reduction: the linalg.fill 0 that seeds the accumulator is the result.
      const uint32_t zero = 0;
      B200_RETURN_IF_ERROR(iree_b200_command_buffer_fill_buffer(
          cb, iree_b200_buffer_ref_t{buffers[2], 0, bytes[2]}, &zero, 4));
    } else if (K > 0) {
      B200_RETURN_IF_ERROR(iree_b200_ops_record_matmul(
          ops, cb, element_type, batch, M, N, K, flags,
          iree_b200_buffer_ref_t{buffers[0], 0, bytes[0]},
          iree_b200_buffer_ref_t{buffers[1], 0, bytes[1]},
          iree_b200_buffer_ref_t{buffers[2], 0, bytes[2]}));
    }
    B200_RETURN_IF_ERROR(iree_b200_command_buffer_end(cb));
    B200_RETURN_IF_ERROR(
        iree_b200_device_queue_execute(device, 0, nullptr, nullptr, 0, nullptr, nullptr, cb));
    B200_RETURN_IF_ERROR(iree_b200_device_transfer_d2h(device, buffers[2], 0, out, bytes[2]));
    return iree_b200_device_wait_idle(device);
  };
  status = run();
  std::string message = status != IREE_B200_STATUS_OK ? iree_b200_status_message() : "";
  if (status != IREE_B200_STATUS_OK) iree_b200_device_wait_idle(device);
  iree_b200_command_buffer_release(cb);
  for (auto* b : buffers) iree_b200_buffer_release(b);
  if (status != IREE_B200_STATUS_OK) return make_status(status, "%s", message.c_str());
  return IREE_B200_STATUS_OK;
}

}  // extern "C"
