// executable.cc — "CDA1" executable blobs → loaded CUDA modules and kernel parameters.
//
// Stands in for runtime/src/iree/hal/drivers/cuda/{nop_executable_cache.c,
// native_executable.c} and hal/utils/executable_header.c. The wire format is the
// reference's (schemas/cuda_executable_def.fbs behind the 64-byte
// iree_flatbuffer_file_header_t, base/internal/flatcc/parsing.h:41-53); the flatbuffer is
// walked by a small bounds-checked reader written for this schema instead of
// flatcc-generated code, and treated as untrusted input.

#include "internal.h"

namespace b200hal {

//===----------------------------------------------------------------------===//
// Header + flatbuffer reader (host only; exercised by the CPU tests)
//===----------------------------------------------------------------------===//

static const char kFileIdentifier[4] = {'C', 'D', 'A', '1'};
static const size_t kHeaderSize = 64;

iree_b200_status_t read_executable_header(const void* data, size_t data_length,
                                          bool unsafe_infer_size, const uint8_t** out_fb,
                                          size_t* out_fb_length) {
  if (!data) return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "executable data is NULL");
  if (!unsafe_infer_size && data_length < kHeaderSize) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                       "executable flatbuffer data is not present or less than header size");
  }
  const uint8_t* bytes = static_cast<const uint8_t*>(data);
  if (memcmp(bytes, kFileIdentifier, 4) != 0) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                       "executable file identifier does not match; expected `CDA1`");
  }
  uint32_t version;
  uint64_t content_size;
  memcpy(&version, bytes + 4, 4);
  memcpy(&content_size, bytes + 8, 8);
  if (content_size == 0) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                       "executable flatbuffer reported as zero-length");
  }
  if (version != 0) {
    return make_status(IREE_B200_STATUS_INCOMPATIBLE,
                       "executable version %u not compatible with this build", version);
  }
  if (!unsafe_infer_size && content_size > data_length - kHeaderSize) {
    return make_status(IREE_B200_STATUS_OUT_OF_RANGE,
                       "executable flatbuffer size prefix out of bounds (size is %llu but only %zu "
                       "bytes available)",
                       (unsigned long long)content_size, data_length - kHeaderSize);
  }
  *out_fb = bytes + kHeaderSize;
  *out_fb_length = (size_t)content_size;
  return IREE_B200_STATUS_OK;
}

namespace {

struct Reader {
  const uint8_t* base;
  size_t size;
  bool ok = true;
  const char* error = "";

  bool in_bounds(size_t pos, size_t len) const { return pos <= size && len <= size - pos; }
  bool fail(const char* what) {
    if (ok) {
      ok = false;
      error = what;
    }
    return false;
  }
  template <typename T>
  T read(size_t pos) {
    T v{};
    if (!in_bounds(pos, sizeof(T))) {
      fail("scalar read out of bounds");
      return v;
    }
    memcpy(&v, base + pos, sizeof(T));
    return v;
  }
};

struct Table {
  size_t pos = 0;      // 0 = absent
  size_t vtable = 0;
  uint16_t vtable_size = 0;
  uint16_t table_size = 0;
};

static Table open_table(Reader& r, size_t pos) {
  Table t;
  if (!r.ok) return t;
  if (pos % 4 != 0 || !r.in_bounds(pos, 4)) {
    r.fail("table position invalid");
    return t;
  }
  const int32_t soff = r.read<int32_t>(pos);
  const int64_t vt = (int64_t)pos - (int64_t)soff;
  if (vt < 0 || vt % 2 != 0 || !r.in_bounds((size_t)vt, 4)) {
    r.fail("vtable position invalid");
    return t;
  }
  t.vtable = (size_t)vt;
  t.vtable_size = r.read<uint16_t>(t.vtable);
  t.table_size = r.read<uint16_t>(t.vtable + 2);
  if (t.vtable_size < 4 || t.vtable_size % 2 != 0 || !r.in_bounds(t.vtable, t.vtable_size) ||
      t.table_size < 4 || !r.in_bounds(pos, t.table_size)) {
    r.fail("vtable/table size invalid");
    return Table();
  }
  t.pos = pos;
  return t;
}

// Position of field |id| inside the table, or 0 when absent.
static size_t field_pos(Reader& r, const Table& t, int id, size_t field_size) {
  if (!r.ok || !t.pos) return 0;
  const size_t slot = 4 + 2 * (size_t)id;
  if (slot + 2 > t.vtable_size) return 0;
  const uint16_t off = r.read<uint16_t>(t.vtable + slot);
  if (off == 0) return 0;
  if ((size_t)off + field_size > t.table_size) {
    r.fail("field exceeds table");
    return 0;
  }
  return t.pos + off;
}

// Follows a uoffset stored at |pos|.
static size_t follow(Reader& r, size_t pos) {
  if (!r.ok || !pos) return 0;
  const uint32_t off = r.read<uint32_t>(pos);
  if (off == 0 || !r.in_bounds(pos, off) || !r.in_bounds(pos + off, 4)) {
    r.fail("offset out of bounds");
    return 0;
  }
  return pos + off;
}

static bool read_string(Reader& r, size_t pos, const uint8_t** out_data, size_t* out_length) {
  *out_data = nullptr;
  *out_length = 0;
  if (!r.ok || !pos) return false;
  const uint32_t len = r.read<uint32_t>(pos);
  if (!r.in_bounds(pos + 4, (size_t)len + 1)) return r.fail("string out of bounds");
  if (r.base[pos + 4 + len] != 0) return r.fail("string not NUL terminated");
  *out_data = r.base + pos + 4;
  *out_length = len;
  return true;
}

// Returns the element count and sets |*out_first| to the first element position.
static uint32_t open_vector(Reader& r, size_t pos, size_t elem_size, size_t* out_first) {
  *out_first = 0;
  if (!r.ok || !pos) return 0;
  const uint32_t count = r.read<uint32_t>(pos);
  if ((uint64_t)count * elem_size > r.size || !r.in_bounds(pos + 4, (size_t)count * elem_size)) {
    r.fail("vector out of bounds");
    return 0;
  }
  *out_first = pos + 4;
  return count;
}

}  // namespace

iree_b200_status_t parse_executable_def(const uint8_t* fb, size_t fb_length,
                                        ExecutableDefView* out_view) {
  out_view->exports.clear();
  out_view->modules.clear();
  if (fb_length < 16) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                       "flatbuffer verification failed: buffer too small");
  }
  Reader r{fb, fb_length};
  if (memcmp(fb + 4, kFileIdentifier, 4) != 0) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                       "flatbuffer verification failed: identifier mismatch");
  }
  const uint32_t root_off = r.read<uint32_t>(0);
  Table root = open_table(r, root_off);

  // modules:[ModuleDef] (field 1)
  {
    size_t first = 0;
    const uint32_t count = open_vector(r, follow(r, field_pos(r, root, 1, 4)), 4, &first);
    for (uint32_t i = 0; r.ok && i < count; ++i) {
      Table module = open_table(r, follow(r, first + 4 * i));
      const uint8_t* data = nullptr;
      size_t length = 0;
      read_string(r, follow(r, field_pos(r, module, 0, 4)), &data, &length);
      if (r.ok && length == 0) {
        return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "modules[%u] contents are empty", i);
      }
      out_view->modules.emplace_back(data, length);
    }
  }

  // exports:[ExportDef] (field 0)
  {
    size_t first = 0;
    const uint32_t count = open_vector(r, follow(r, field_pos(r, root, 0, 4)), 4, &first);
    for (uint32_t i = 0; r.ok && i < count; ++i) {
      ExportDefView e;
      Table t = open_table(r, follow(r, first + 4 * i));
      if (size_t p = field_pos(r, t, 0, 4)) e.module_ordinal = r.read<uint32_t>(p);
      const uint8_t* name = nullptr;
      size_t name_length = 0;
      read_string(r, follow(r, field_pos(r, t, 1, 4)), &name, &name_length);
      if (name) e.kernel_name.assign(reinterpret_cast<const char*>(name), name_length);
      if (size_t p = field_pos(r, t, 2, 12)) {
        e.has_block_dims = true;
        for (int d = 0; d < 3; ++d) e.block_dims[d] = r.read<uint32_t>(p + 4 * d);
      }
      if (size_t p = field_pos(r, t, 3, 4)) e.block_shared_memory_size = r.read<uint32_t>(p);
      if (size_t p = field_pos(r, t, 4, 4)) e.constant_count = r.read<uint32_t>(p);
      size_t flags_first = 0;
      const uint32_t flag_count =
          open_vector(r, follow(r, field_pos(r, t, 5, 4)), 8, &flags_first);
      for (uint32_t j = 0; r.ok && j < flag_count; ++j) {
        e.binding_flags.push_back(r.read<uint64_t>(flags_first + 8 * j));
      }
      out_view->exports.push_back(std::move(e));
    }
  }

  if (!r.ok) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "flatbuffer verification failed: %s",
                       r.error);
  }

  // Semantic checks that need no device (native_executable.c:150-195,223-245).
  for (size_t i = 0; i < out_view->exports.size(); ++i) {
    const ExportDefView& e = out_view->exports[i];
    if (e.module_ordinal >= out_view->modules.size()) {
      return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                         "exports[%zu] module_ordinal %u is out of bounds %zu", i, e.module_ordinal,
                         out_view->modules.size());
    }
    if (e.kernel_name.empty()) {
      return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "exports[%zu] name is empty", i);
    }
    if (!e.has_block_dims) {
      return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "exports[%zu] blocks dims are missing",
                         i);
    }
    if (e.constant_count > IREE_B200_MAX_DISPATCH_CONSTANT_COUNT) {
      return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                         "exports[%zu] constant_count %u exceeds maximum of %u", i,
                         e.constant_count, IREE_B200_MAX_DISPATCH_CONSTANT_COUNT);
    }
    if (e.binding_flags.size() > IREE_B200_MAX_DISPATCH_BINDING_COUNT) {
      return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                         "exports[%zu] binding_flags count %zu exceeds maximum of %u", i,
                         e.binding_flags.size(), IREE_B200_MAX_DISPATCH_BINDING_COUNT);
    }
  }
  return IREE_B200_STATUS_OK;
}

//===----------------------------------------------------------------------===//
// Tensor maps
//===----------------------------------------------------------------------===//

iree_b200_status_t encode_tensor_map(const CudaSymbols* cu, const iree_b200_tma_map_t& map,
                                     uint64_t base_pointer, const uint32_t* constants,
                                     size_t constant_count, CUtensorMap* out_map) {
  if (map.rank < 1 || map.rank > IREE_B200_TMA_MAX_RANK) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "TMA recipe rank %u unsupported",
                       map.rank);
  }
  cuuint64_t dims[IREE_B200_TMA_MAX_RANK];
  cuuint64_t strides[IREE_B200_TMA_MAX_RANK];       // globalStrides[i] = stride of dim i + 1
  cuuint32_t box[IREE_B200_TMA_MAX_RANK];
  cuuint32_t elem_strides[IREE_B200_TMA_MAX_RANK];
  auto resolve = [&](uint32_t v, uint64_t* out) -> bool {
    if (v & 0x80000000u) {
      const uint32_t index = v & 0x7fffffffu;
      if (index >= constant_count) return false;
      v = constants[index];
    }
    *out = v;
    return true;
  };
  uint32_t order[IREE_B200_TMA_MAX_RANK];
  for (uint32_t d = 0; d < map.rank; ++d) {
    order[d] = map.order ? (map.order >> (4 * d)) & 0xfu : d;
    uint64_t extent = 0, tile = 0;
    if (!resolve(map.dim[d], &extent) || !resolve(map.box[d], &tile)) {
      return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                         "TMA recipe refers to a push constant beyond the %zu provided",
                         constant_count);
    }
    if (extent == 0 || tile == 0 || tile > 256) {
      return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                         "TMA tensor extent / box of dim %u is zero or too large", d);
    }
    dims[d] = extent;
    box[d] = (cuuint32_t)tile;
    elem_strides[d] = 1;
  }
  if (order[0] != 0) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                       "TMA recipe: map dimension 0 must be the dense-innermost one");
  }
  for (uint32_t d = 0; d < map.rank; ++d) {
    uint64_t stride = map.elem_size;
    for (uint32_t e = 0; e < map.rank; ++e) {
      if (order[e] < order[d]) stride *= dims[e];
    }
    if (d > 0) strides[d - 1] = stride;
  }
  if (base_pointer % 16 != 0) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                       "TMA-fed binding %u is not 16-byte aligned (0x%llx)", map.binding,
                       (unsigned long long)base_pointer);
  }
  for (uint32_t d = 0; d + 1 < map.rank; ++d) {
    if (strides[d] % 16 != 0) {
      return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                         "TMA-fed binding %u: byte stride %llu of dim %u is not a multiple of 16",
                         map.binding, (unsigned long long)strides[d], d + 1);
    }
  }
  uint64_t swizzle = 0;
  if (!resolve(map.swizzle, &swizzle) || swizzle > IREE_B200_TMA_SWIZZLE_128B) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "TMA recipe: bad swizzle mode");
  }
  // Measured on the B200 (tools/tma4d_probe.cu): a swizzled box whose inner row is narrower
  // than the swizzle span faults with an illegal address; refuse it here instead.
  if (swizzle && (uint64_t)box[0] * map.elem_size < (16ull << swizzle)) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                       "TMA recipe: inner box of %llu bytes is narrower than the %llu-byte swizzle",
                       (unsigned long long)box[0] * map.elem_size, 16ull << swizzle);
  }
  B200_CU(cu, cuTensorMapEncodeTiled(
                  out_map, (CUtensorMapDataType)map.elem_type, map.rank,
                  reinterpret_cast<void*>(base_pointer), dims, strides, box, elem_strides,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, (CUtensorMapSwizzle)swizzle,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE));
  return IREE_B200_STATUS_OK;
}

}  // namespace b200hal

using namespace b200hal;

extern "C" {

iree_b200_status_t iree_b200_executable_cache_create(iree_b200_device_t* device,
                                                     const char* identifier,
                                                     iree_b200_executable_cache_t** out_cache) {
  B200_REQUIRE_ARG(device && out_cache);
  auto* cache = new iree_b200_executable_cache_t();
  cache->device = device;
  cache->identifier = identifier ? identifier : "";
  iree_b200_device_retain(device);
  *out_cache = cache;
  return IREE_B200_STATUS_OK;
}

void iree_b200_executable_cache_release(iree_b200_executable_cache_t* cache) {
  if (!cache || cache->ref_count.fetch_sub(1) != 1) return;
  iree_b200_device_release(cache->device);
  delete cache;
}

int32_t iree_b200_executable_cache_can_prepare_format(iree_b200_executable_cache_t* cache,
                                                      const char* executable_format) {
  (void)cache;
  return executable_format && strcmp(executable_format, "PTXE") == 0;
}

iree_b200_status_t iree_b200_executable_infer_format(const void* executable_data,
                                                     size_t data_length, size_t format_capacity,
                                                     char* out_executable_format,
                                                     size_t* out_inferred_size) {
  B200_REQUIRE_ARG(out_executable_format && out_inferred_size);
  const uint8_t* fb = nullptr;
  size_t fb_length = 0;
  B200_RETURN_IF_ERROR(
      read_executable_header(executable_data, data_length, data_length == 0, &fb, &fb_length));
  ExecutableDefView view;
  B200_RETURN_IF_ERROR(parse_executable_def(fb, fb_length, &view));
  if (format_capacity <= 4) {
    return make_status(IREE_B200_STATUS_OUT_OF_RANGE, "executable format buffer too small");
  }
  memcpy(out_executable_format, "PTXE", 5);
  *out_inferred_size = 64 + fb_length;
  return IREE_B200_STATUS_OK;
}

static iree_b200_status_t executable_load(iree_b200_executable_t* exe, const ExecutableDefView& view) {
  iree_b200_device_t* device = exe->device;
  const CudaSymbols* cu = device->cu;
  B200_CU(cu, cuCtxSetCurrent(device->cu_context));

  // Device-limit checks (native_executable.c:192-221).
  for (size_t i = 0; i < view.exports.size(); ++i) {
    const ExportDefView& e = view.exports[i];
    for (int d = 0; d < 3; ++d) {
      if (e.block_dims[d] > (uint32_t)device->max_block_dims[d]) {
        return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                           "exports[%zu] block dims %ux%ux%u exceeds device maximum %dx%dx%d", i,
                           e.block_dims[0], e.block_dims[1], e.block_dims[2],
                           device->max_block_dims[0], device->max_block_dims[1],
                           device->max_block_dims[2]);
      }
    }
    if (e.block_shared_memory_size > (uint32_t)device->max_smem_optin) {
      return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                         "exports[%zu] requires %uB of shared memory and exceeds the device maximum "
                         "of %dB per block",
                         i, e.block_shared_memory_size, device->max_smem_optin);
    }
  }

  for (size_t i = 0; i < view.modules.size(); ++i) {
    // cuModuleLoadDataEx with an error log (native_executable.c:360-393). The image may
    // be PTX text or a cubin/fatbin; the driver sniffs the container.
    char error_log[8192] = {0};
    CUjit_option options[] = {CU_JIT_ERROR_LOG_BUFFER, CU_JIT_ERROR_LOG_BUFFER_SIZE_BYTES};
    void* values[] = {error_log, reinterpret_cast<void*>(sizeof(error_log))};
    CUmodule module = nullptr;
    // PTX must be NUL-terminated: flatbuffer strings are. ELF images carry their size.
    // Flatbuffer strings are only 4-byte aligned; give the driver an 8-byte aligned copy.
    std::vector<uint64_t> image((view.modules[i].second + 1 + 7) / 8, 0);
    memcpy(image.data(), view.modules[i].first, view.modules[i].second);
    CUresult r = cu->cuModuleLoadDataEx(&module, image.data(), 2, options, values);
    if (r != CUDA_SUCCESS) {
      iree_b200_status_t status = cu_status(cu, r, "cuModuleLoadDataEx", __FILE__, __LINE__);
      std::string message = iree_b200_status_message();
      return make_status(status, "%s; modules[%zu] log: %s", message.c_str(), i, error_log);
    }
    exe->modules.push_back(module);
  }

  exe->exports.resize(view.exports.size());
  for (size_t i = 0; i < view.exports.size(); ++i) {
    const ExportDefView& e = view.exports[i];
    KernelParams& k = exe->exports[i];
    k.name = e.kernel_name;
    CUmodule module = exe->modules[e.module_ordinal];
    CUresult r = cu->cuModuleGetFunction(&k.function, module, k.name.c_str());
    if (r != CUDA_SUCCESS || !k.function) {
      iree_b200_status_t status = cu_status(cu, r, "cuModuleGetFunction", __FILE__, __LINE__);
      std::string message = iree_b200_status_message();
      return make_status(status == IREE_B200_STATUS_OK ? IREE_B200_STATUS_NOT_FOUND : status,
                         "exports[%zu] kernel `%s` not found in modules[%u]: %s", i, k.name.c_str(),
                         e.module_ordinal, message.c_str());
    }
    for (int d = 0; d < 3; ++d) k.block_dims[d] = e.block_dims[d];
    k.block_shared_memory_size = e.block_shared_memory_size;
    k.constant_count = e.constant_count;
    k.binding_count = (uint32_t)e.binding_flags.size();
    for (size_t j = 0; j < e.binding_flags.size(); ++j) k.binding_flags[j] = e.binding_flags[j];
    if (k.block_shared_memory_size > 0) {
      B200_CU(cu, cuFuncSetAttribute(k.function, CU_FUNC_ATTRIBUTE_MAX_DYNAMIC_SHARED_SIZE_BYTES,
                                     (int)k.block_shared_memory_size));
    }
    // Optional TMA recipe global shipped beside the kernel (iree_b200_tma_recipe.h); the
    // lookup is the mechanism of lookup_global_by_name (native_executable.c:579-649).
    CUdeviceptr recipe_ptr = 0;
    size_t recipe_size = 0;
    const std::string recipe_name = k.name + "__tma";
    r = cu->cuModuleGetGlobal(&recipe_ptr, &recipe_size, module, recipe_name.c_str());
    if (r == CUDA_SUCCESS) {
      if (recipe_size != sizeof(iree_b200_tma_recipe_t)) {
        return make_status(IREE_B200_STATUS_INCOMPATIBLE,
                           "`%s` is %zu bytes, expected %zu (recipe layout mismatch)",
                           recipe_name.c_str(), recipe_size, sizeof(iree_b200_tma_recipe_t));
      }
      B200_CU(cu, cuMemcpyDtoHAsync(&k.recipe, recipe_ptr, sizeof(k.recipe), device->dispatch_stream));
      B200_CU(cu, cuStreamSynchronize(device->dispatch_stream));
      if (k.recipe.magic != IREE_B200_TMA_RECIPE_MAGIC ||
          k.recipe.map_count > IREE_B200_TMA_MAX_MAPS) {
        return make_status(IREE_B200_STATUS_INCOMPATIBLE, "`%s` has a bad magic or map count",
                           recipe_name.c_str());
      }
      for (uint32_t m = 0; m < k.recipe.map_count; ++m) {
        if (k.recipe.maps[m].binding >= k.binding_count) {
          return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                             "`%s` map %u refers to binding %u of %u", recipe_name.c_str(), m,
                             k.recipe.maps[m].binding, k.binding_count);
        }
      }
      k.has_recipe = true;
    } else if (r != CUDA_ERROR_NOT_FOUND) {
      return cu_status(cu, r, "cuModuleGetGlobal", __FILE__, __LINE__);
    }
  }
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t iree_b200_executable_cache_prepare_executable(
    iree_b200_executable_cache_t* cache, const char* executable_format,
    const void* executable_data, size_t data_length, iree_b200_executable_t** out_executable) {
  B200_REQUIRE_ARG(cache && out_executable);
  *out_executable = nullptr;
  if (!iree_b200_executable_cache_can_prepare_format(cache, executable_format)) {
    return make_status(IREE_B200_STATUS_INCOMPATIBLE, "executable format `%s` is not supported",
                       executable_format ? executable_format : "(null)");
  }
  const uint8_t* fb = nullptr;
  size_t fb_length = 0;
  B200_RETURN_IF_ERROR(read_executable_header(executable_data, data_length,
                                              /*unsafe_infer_size=*/false, &fb, &fb_length));
  ExecutableDefView view;
  B200_RETURN_IF_ERROR(parse_executable_def(fb, fb_length, &view));

  auto* exe = new iree_b200_executable_t();
  exe->device = cache->device;
  iree_b200_device_retain(exe->device);
  iree_b200_status_t status = executable_load(exe, view);
  if (status != IREE_B200_STATUS_OK) {
    std::string message = iree_b200_status_message();
    iree_b200_executable_release(exe);
    return make_status(status, "%s", message.c_str());
  }
  *out_executable = exe;
  return IREE_B200_STATUS_OK;
}

void iree_b200_executable_retain(iree_b200_executable_t* executable) {
  if (executable) executable->ref_count.fetch_add(1);
}

void iree_b200_executable_release(iree_b200_executable_t* executable) {
  if (!executable || executable->ref_count.fetch_sub(1) != 1) return;
  iree_b200_device_t* device = executable->device;
  if (device) {  // NULL for the module-less executable of an offline ops object
    device->cu->cuCtxSetCurrent(device->cu_context);
    for (CUmodule module : executable->modules) device->cu->cuModuleUnload(module);
  }
  delete executable;
  iree_b200_device_release(device);
}

int32_t iree_b200_executable_export_count(const iree_b200_executable_t* executable) {
  return executable ? (int32_t)executable->exports.size() : 0;
}

iree_b200_status_t iree_b200_executable_export_info(const iree_b200_executable_t* executable,
                                                    int32_t export_ordinal,
                                                    iree_b200_export_info_t* out_info) {
  B200_REQUIRE_ARG(executable && out_info);
  if (export_ordinal < 0 || (size_t)export_ordinal >= executable->exports.size()) {
    return make_status(IREE_B200_STATUS_OUT_OF_RANGE, "export ordinal %d out of range; executable has %zu exports",
                       export_ordinal, executable->exports.size());
  }
  const KernelParams& k = executable->exports[export_ordinal];
  memset(out_info, 0, sizeof(*out_info));
  out_info->kernel_name = k.name.c_str();
  for (int d = 0; d < 3; ++d) out_info->block_dims[d] = k.block_dims[d];
  out_info->block_shared_memory_size = k.block_shared_memory_size;
  out_info->constant_count = k.constant_count;
  out_info->binding_count = k.binding_count;
  out_info->tma_map_count = k.has_recipe ? k.recipe.map_count : 0;
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t iree_b200_executable_lookup_export_by_name(
    const iree_b200_executable_t* executable, const char* name, int32_t* out_export_ordinal) {
  B200_REQUIRE_ARG(executable && name && out_export_ordinal);
  for (size_t i = 0; i < executable->exports.size(); ++i) {
    if (executable->exports[i].name == name) {
      *out_export_ordinal = (int32_t)i;
      return IREE_B200_STATUS_OK;
    }
  }
  *out_export_ordinal = -1;
  return make_status(IREE_B200_STATUS_NOT_FOUND, "no export named `%s`", name);
}

}  // extern "C"
