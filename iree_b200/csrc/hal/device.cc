// device.cc — status, dynamic symbols, device, buffers, semaphores, queue submission.
//
// Stands in for the corresponding pieces of runtime/src/iree/hal/drivers/cuda
// (cuda_dynamic_symbols.c, cuda_status_util.h, cuda_device.c, cuda_allocator.c,
// cuda_buffer.c, event_semaphore.c); see include/iree_b200_hal.h for the per-entry
// citations. One primary context and one dispatch stream per device, as in the
// reference (cuda_device.c:573-587).

#include <dlfcn.h>

#include <chrono>
#include <thread>

#include "internal.h"

namespace b200hal {

static thread_local char tls_message[1024] = "";

iree_b200_status_t make_status(iree_b200_status_t code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(tls_message, sizeof(tls_message), fmt, ap);
  va_end(ap);
  return code;
}

#define B200_STR_(x) #x
#define B200_STR(x) B200_STR_(x)

static std::once_flag cuda_once;
static CudaSymbols cuda_table;
static iree_b200_status_t cuda_load_status = IREE_B200_STATUS_UNAVAILABLE;
static std::string cuda_load_message = "not loaded";

static void load_cuda() {
  void* lib = dlopen("libcuda.so.1", RTLD_NOW | RTLD_GLOBAL);
  if (!lib) lib = dlopen("libcuda.so", RTLD_NOW | RTLD_GLOBAL);
  if (!lib) {
    cuda_load_message = std::string("CUDA driver library not found: ") + dlerror();
    return;
  }
#define B200_LOAD(name)                                                         \
  cuda_table.name = reinterpret_cast<decltype(cuda_table.name)>(dlsym(lib, B200_STR(name))); \
  if (!cuda_table.name) {                                                       \
    cuda_load_message = "CUDA driver symbol missing: " B200_STR(name);          \
    return;                                                                     \
  }
  B200_CUDA_SYMBOLS(B200_LOAD)
#undef B200_LOAD
  CUresult r = cuda_table.cuInit(0);
  if (r != CUDA_SUCCESS) {
    const char* name = nullptr;
    cuda_table.cuGetErrorName(r, &name);
    cuda_load_message = std::string("cuInit failed: ") + (name ? name : "?");
    return;
  }
  cuda_load_status = IREE_B200_STATUS_OK;
}

iree_b200_status_t cuda_symbols(const CudaSymbols** out_syms) {
  std::call_once(cuda_once, load_cuda);
  if (cuda_load_status != IREE_B200_STATUS_OK) {
    return make_status(cuda_load_status, "%s", cuda_load_message.c_str());
  }
  *out_syms = &cuda_table;
  return IREE_B200_STATUS_OK;
}

static std::once_flag nccl_once;
static NcclSymbols nccl_table;
static iree_b200_status_t nccl_load_status = IREE_B200_STATUS_UNAVAILABLE;
static std::string nccl_load_message = "not loaded";

static void load_nccl() {
  // RTLD_NOLOAD first: inside a process that already carries an NCCL (e.g. the one
  // bundled with torch) reuse it rather than loading a second copy.
  void* lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
  if (!lib) lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  if (!lib) {
    nccl_load_message = std::string("NCCL library not found: ") + dlerror();
    return;
  }
#define B200_LOAD(name)                                                         \
  nccl_table.name = reinterpret_cast<decltype(nccl_table.name)>(dlsym(lib, #name)); \
  if (!nccl_table.name) {                                                       \
    nccl_load_message = "NCCL symbol missing: " #name;                          \
    return;                                                                     \
  }
  B200_NCCL_SYMBOLS(B200_LOAD)
#undef B200_LOAD
  nccl_load_status = IREE_B200_STATUS_OK;
}

iree_b200_status_t nccl_symbols(const NcclSymbols** out_syms) {
  std::call_once(nccl_once, load_nccl);
  if (nccl_load_status != IREE_B200_STATUS_OK) {
    return make_status(nccl_load_status, "%s", nccl_load_message.c_str());
  }
  *out_syms = &nccl_table;
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t cu_status(const CudaSymbols* syms, CUresult result, const char* expr,
                             const char* file, int line) {
  const char* name = nullptr;
  const char* desc = nullptr;
  syms->cuGetErrorName(result, &name);
  syms->cuGetErrorString(result, &desc);
  iree_b200_status_t code = IREE_B200_STATUS_INTERNAL;
  switch (result) {
    case CUDA_ERROR_INVALID_VALUE:
      code = IREE_B200_STATUS_INVALID_ARGUMENT;
      break;
    case CUDA_ERROR_OUT_OF_MEMORY:
      code = IREE_B200_STATUS_RESOURCE_EXHAUSTED;
      break;
    case CUDA_ERROR_NOT_FOUND:
      code = IREE_B200_STATUS_NOT_FOUND;
      break;
    case CUDA_ERROR_NO_DEVICE:
    case CUDA_ERROR_INVALID_DEVICE:
      code = IREE_B200_STATUS_UNAVAILABLE;
      break;
    case CUDA_ERROR_NOT_SUPPORTED:
      code = IREE_B200_STATUS_UNIMPLEMENTED;
      break;
    default:
      break;
  }
  const char* base = strrchr(file, '/');
  return make_status(code, "%s:%d: CUDA driver error %s (%d): %s; in `%s`", base ? base + 1 : file,
                     line, name ? name : "?", (int)result, desc ? desc : "?", expr);
}

iree_b200_status_t nccl_status(const NcclSymbols* syms, ncclResult_t result, const char* expr,
                               const char* file, int line) {
  const char* base = strrchr(file, '/');
  return make_status(IREE_B200_STATUS_INTERNAL, "%s:%d: NCCL error %d: %s (%s); in `%s`",
                     base ? base + 1 : file, line, (int)result, syms->ncclGetErrorString(result),
                     syms->ncclGetLastError(nullptr), expr);
}

}  // namespace b200hal

using namespace b200hal;

extern "C" {

const char* iree_b200_status_message(void) { return tls_message; }

const char* iree_b200_status_code_string(iree_b200_status_t status) {
  switch (status) {
    case IREE_B200_STATUS_OK: return "OK";
    case IREE_B200_STATUS_CANCELLED: return "CANCELLED";
    case IREE_B200_STATUS_UNKNOWN: return "UNKNOWN";
    case IREE_B200_STATUS_INVALID_ARGUMENT: return "INVALID_ARGUMENT";
    case IREE_B200_STATUS_DEADLINE_EXCEEDED: return "DEADLINE_EXCEEDED";
    case IREE_B200_STATUS_NOT_FOUND: return "NOT_FOUND";
    case IREE_B200_STATUS_ALREADY_EXISTS: return "ALREADY_EXISTS";
    case IREE_B200_STATUS_PERMISSION_DENIED: return "PERMISSION_DENIED";
    case IREE_B200_STATUS_RESOURCE_EXHAUSTED: return "RESOURCE_EXHAUSTED";
    case IREE_B200_STATUS_FAILED_PRECONDITION: return "FAILED_PRECONDITION";
    case IREE_B200_STATUS_ABORTED: return "ABORTED";
    case IREE_B200_STATUS_OUT_OF_RANGE: return "OUT_OF_RANGE";
    case IREE_B200_STATUS_UNIMPLEMENTED: return "UNIMPLEMENTED";
    case IREE_B200_STATUS_INTERNAL: return "INTERNAL";
    case IREE_B200_STATUS_UNAVAILABLE: return "UNAVAILABLE";
    case IREE_B200_STATUS_DATA_LOSS: return "DATA_LOSS";
    case IREE_B200_STATUS_INCOMPATIBLE: return "INCOMPATIBLE";
    default: return "?";
  }
}

//===----------------------------------------------------------------------===//
// Device
//===----------------------------------------------------------------------===//

void iree_b200_device_params_initialize(iree_b200_device_params_t* out_params) {
  if (!out_params) return;
  memset(out_params, 0, sizeof(*out_params));
  out_params->command_buffer_mode = IREE_B200_COMMAND_BUFFER_MODE_GRAPH;
  out_params->overlap_collectives = 0;
}

static iree_b200_status_t device_initialize(iree_b200_device_t* d) {
  const CudaSymbols* cu = d->cu;
  int count = 0;
  B200_CU(cu, cuDeviceGetCount(&count));
  if (d->ordinal < 0 || d->ordinal >= count) {
    return make_status(IREE_B200_STATUS_NOT_FOUND, "CUDA device ordinal %d out of range (%d devices)",
                       d->ordinal, count);
  }
  B200_CU(cu, cuDeviceGet(&d->cu_device, d->ordinal));
  B200_CU(cu, cuDeviceGetAttribute(&d->cc_major, CU_DEVICE_ATTRIBUTE_COMPUTE_CAPABILITY_MAJOR,
                                   d->cu_device));
  B200_CU(cu, cuDeviceGetAttribute(&d->cc_minor, CU_DEVICE_ATTRIBUTE_COMPUTE_CAPABILITY_MINOR,
                                   d->cu_device));
  B200_CU(cu, cuDeviceGetAttribute(&d->sm_count, CU_DEVICE_ATTRIBUTE_MULTIPROCESSOR_COUNT,
                                   d->cu_device));
  B200_CU(cu, cuDeviceGetAttribute(&d->max_smem_optin,
                                   CU_DEVICE_ATTRIBUTE_MAX_SHARED_MEMORY_PER_BLOCK_OPTIN,
                                   d->cu_device));
  B200_CU(cu, cuDeviceGetAttribute(&d->max_block_dims[0], CU_DEVICE_ATTRIBUTE_MAX_BLOCK_DIM_X,
                                   d->cu_device));
  B200_CU(cu, cuDeviceGetAttribute(&d->max_block_dims[1], CU_DEVICE_ATTRIBUTE_MAX_BLOCK_DIM_Y,
                                   d->cu_device));
  B200_CU(cu, cuDeviceGetAttribute(&d->max_block_dims[2], CU_DEVICE_ATTRIBUTE_MAX_BLOCK_DIM_Z,
                                   d->cu_device));
  B200_CU(cu, cuDeviceGetAttribute(&d->max_threads_per_block,
                                   CU_DEVICE_ATTRIBUTE_MAX_THREADS_PER_BLOCK, d->cu_device));
  B200_CU(cu, cuDevicePrimaryCtxRetain(&d->cu_context, d->cu_device));
  B200_CU(cu, cuCtxSetCurrent(d->cu_context));
  if (d->params.external_dispatch_stream) {
    d->dispatch_stream = reinterpret_cast<CUstream>(d->params.external_dispatch_stream);
    d->owns_dispatch_stream = false;
  } else {
    B200_CU(cu, cuStreamCreate(&d->dispatch_stream, CU_STREAM_NON_BLOCKING));
    d->owns_dispatch_stream = true;
  }
  if (d->params.overlap_collectives) {
    int least = 0, greatest = 0;
    B200_CU(cu, cuCtxGetStreamPriorityRange(&least, &greatest));
    B200_CU(cu, cuStreamCreateWithPriority(&d->collective_stream, CU_STREAM_NON_BLOCKING, greatest));
  }
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t iree_b200_device_create(int32_t ordinal, const iree_b200_device_params_t* params,
                                           iree_b200_device_t** out_device) {
  B200_REQUIRE_ARG(out_device);
  *out_device = nullptr;
  const CudaSymbols* cu = nullptr;
  B200_RETURN_IF_ERROR(cuda_symbols(&cu));
  auto* d = new iree_b200_device_t();
  d->cu = cu;
  d->ordinal = ordinal;
  if (params) {
    d->params = *params;
  } else {
    iree_b200_device_params_initialize(&d->params);
  }
  iree_b200_status_t status = device_initialize(d);
  if (status != IREE_B200_STATUS_OK) {
    iree_b200_device_release(d);
    return status;
  }
  *out_device = d;
  return IREE_B200_STATUS_OK;
}

void iree_b200_device_retain(iree_b200_device_t* device) {
  if (device) device->ref_count.fetch_add(1);
}

void iree_b200_device_release(iree_b200_device_t* device) {
  if (!device || device->ref_count.fetch_sub(1) != 1) return;
  const CudaSymbols* cu = device->cu;
  if (device->cu_context) {
    cu->cuCtxSetCurrent(device->cu_context);
    if (device->dispatch_stream && device->owns_dispatch_stream) {
      cu->cuStreamSynchronize(device->dispatch_stream);
      cu->cuStreamDestroy(device->dispatch_stream);
    }
    if (device->collective_stream) {
      cu->cuStreamSynchronize(device->collective_stream);
      cu->cuStreamDestroy(device->collective_stream);
    }
    cu->cuDevicePrimaryCtxRelease(device->cu_device);
  }
  delete device;
}

iree_b200_status_t iree_b200_device_query_i64(iree_b200_device_t* device, const char* category,
                                              const char* key, int64_t* out_value) {
  B200_REQUIRE_ARG(device && category && key && out_value);
  *out_value = 0;
  if (!strcmp(category, "hal.device.id")) {
    *out_value = !strcmp(key, "cuda") || !strcmp(key, "cuda*") ? 1 : 0;
    return IREE_B200_STATUS_OK;
  }
  if (!strcmp(category, "hal.executable.format")) {
    *out_value = !strcmp(key, "cuda-nvptx-fb") ? 1 : 0;
    return IREE_B200_STATUS_OK;
  }
  if (!strcmp(category, "cuda.device")) {
    if (!strcmp(key, "compute_capability_major")) {
      *out_value = device->cc_major;
      return IREE_B200_STATUS_OK;
    }
    if (!strcmp(key, "compute_capability_minor")) {
      *out_value = device->cc_minor;
      return IREE_B200_STATUS_OK;
    }
  }
  if (!strcmp(category, "b200.device")) {
    if (!strcmp(key, "multiprocessor_count")) {
      *out_value = device->sm_count;
      return IREE_B200_STATUS_OK;
    }
    if (!strcmp(key, "max_shared_memory_per_block_optin")) {
      *out_value = device->max_smem_optin;
      return IREE_B200_STATUS_OK;
    }
  }
  return make_status(IREE_B200_STATUS_NOT_FOUND, "unknown device configuration key value '%s :: %s'",
                     category, key);
}

uint64_t iree_b200_device_dispatch_stream(iree_b200_device_t* device) {
  return device ? reinterpret_cast<uint64_t>(device->dispatch_stream) : 0;
}

iree_b200_status_t iree_b200_device_wait_idle(iree_b200_device_t* device) {
  B200_REQUIRE_ARG(device);
  const CudaSymbols* cu = device->cu;
  B200_CU(cu, cuCtxSetCurrent(device->cu_context));
  B200_CU(cu, cuStreamSynchronize(device->dispatch_stream));
  if (device->collective_stream) B200_CU(cu, cuStreamSynchronize(device->collective_stream));
  b200hal::reap_in_flight(device, /*everything=*/true);
  return IREE_B200_STATUS_OK;
}

//===----------------------------------------------------------------------===//
// Buffers
//===----------------------------------------------------------------------===//

iree_b200_status_t iree_b200_allocator_allocate_buffer(iree_b200_device_t* device,
                                                       int32_t memory_type, uint64_t byte_length,
                                                       iree_b200_buffer_t** out_buffer) {
  B200_REQUIRE_ARG(device && out_buffer);
  *out_buffer = nullptr;
  const CudaSymbols* cu = device->cu;
  B200_CU(cu, cuCtxSetCurrent(device->cu_context));
  auto* b = new iree_b200_buffer_t();
  b->device = device;
  b->memory_type = memory_type;
  b->byte_length = byte_length;
  const size_t alloc_size = byte_length ? (size_t)byte_length : 1;  // zero-length buffers are legal
  CUresult r;
  if (memory_type == IREE_B200_MEMORY_TYPE_DEVICE_LOCAL) {
    r = cu->cuMemAlloc(&b->device_ptr, alloc_size);
  } else if (memory_type == IREE_B200_MEMORY_TYPE_HOST_LOCAL) {
    r = cu->cuMemHostAlloc(&b->host_ptr, alloc_size, CU_MEMHOSTALLOC_DEVICEMAP);
    if (r == CUDA_SUCCESS) r = cu->cuMemHostGetDevicePointer(&b->device_ptr, b->host_ptr, 0);
  } else {
    delete b;
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "unsupported memory type %d", memory_type);
  }
  if (r != CUDA_SUCCESS) {
    delete b;
    return cu_status(cu, r, "cuMemAlloc/cuMemHostAlloc", __FILE__, __LINE__);
  }
  iree_b200_device_retain(device);
  *out_buffer = b;
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t iree_b200_allocator_import_buffer(iree_b200_device_t* device,
                                                     uint64_t device_pointer, uint64_t byte_length,
                                                     iree_b200_buffer_t** out_buffer) {
  B200_REQUIRE_ARG(device && out_buffer);
  if (!device_pointer && byte_length) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "cannot import a NULL device pointer");
  }
  auto* b = new iree_b200_buffer_t();
  b->device = device;
  b->memory_type = IREE_B200_MEMORY_TYPE_DEVICE_LOCAL;
  b->imported = true;
  b->device_ptr = (CUdeviceptr)device_pointer;
  b->byte_length = byte_length;
  iree_b200_device_retain(device);
  *out_buffer = b;
  return IREE_B200_STATUS_OK;
}

void iree_b200_buffer_retain(iree_b200_buffer_t* buffer) {
  if (buffer) buffer->ref_count.fetch_add(1);
}

void iree_b200_buffer_release(iree_b200_buffer_t* buffer) {
  if (!buffer || buffer->ref_count.fetch_sub(1) != 1) return;
  iree_b200_device_t* device = buffer->device;
  if (!buffer->imported) {
    device->cu->cuCtxSetCurrent(device->cu_context);
    if (buffer->host_ptr) {
      device->cu->cuMemFreeHost(buffer->host_ptr);
    } else if (buffer->device_ptr) {
      device->cu->cuMemFree(buffer->device_ptr);
    }
  }
  delete buffer;
  iree_b200_device_release(device);
}

uint64_t iree_b200_buffer_device_pointer(const iree_b200_buffer_t* buffer) {
  return buffer ? (uint64_t)buffer->device_ptr : 0;
}
void* iree_b200_buffer_host_pointer(const iree_b200_buffer_t* buffer) {
  return buffer ? buffer->host_ptr : nullptr;
}
uint64_t iree_b200_buffer_byte_length(const iree_b200_buffer_t* buffer) {
  return buffer ? buffer->byte_length : 0;
}

static iree_b200_status_t check_range(const iree_b200_buffer_t* buffer, uint64_t offset,
                                      uint64_t length) {
  if (offset > buffer->byte_length || length > buffer->byte_length - offset) {
    return make_status(IREE_B200_STATUS_OUT_OF_RANGE,
                       "range [%llu, +%llu) exceeds buffer of %llu bytes", (unsigned long long)offset,
                       (unsigned long long)length, (unsigned long long)buffer->byte_length);
  }
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t iree_b200_device_transfer_h2d(iree_b200_device_t* device, const void* host,
                                                 iree_b200_buffer_t* target, uint64_t target_offset,
                                                 uint64_t byte_length) {
  B200_REQUIRE_ARG(device && target && (host || !byte_length));
  B200_RETURN_IF_ERROR(check_range(target, target_offset, byte_length));
  if (!byte_length) return IREE_B200_STATUS_OK;
  const CudaSymbols* cu = device->cu;
  std::lock_guard<std::mutex> lock(device->mutex);
  B200_CU(cu, cuCtxSetCurrent(device->cu_context));
  B200_CU(cu, cuMemcpyHtoDAsync(target->device_ptr + target_offset, host, byte_length,
                                device->dispatch_stream));
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t iree_b200_device_transfer_d2h(iree_b200_device_t* device,
                                                 iree_b200_buffer_t* source, uint64_t source_offset,
                                                 void* host, uint64_t byte_length) {
  B200_REQUIRE_ARG(device && source && (host || !byte_length));
  B200_RETURN_IF_ERROR(check_range(source, source_offset, byte_length));
  if (!byte_length) return IREE_B200_STATUS_OK;
  const CudaSymbols* cu = device->cu;
  std::lock_guard<std::mutex> lock(device->mutex);
  B200_CU(cu, cuCtxSetCurrent(device->cu_context));
  B200_CU(cu, cuMemcpyDtoHAsync(host, source->device_ptr + source_offset, byte_length,
                                device->dispatch_stream));
  return IREE_B200_STATUS_OK;
}

//===----------------------------------------------------------------------===//
// Semaphores
//===----------------------------------------------------------------------===//

iree_b200_status_t iree_b200_semaphore_create(iree_b200_device_t* device, uint64_t initial_value,
                                              iree_b200_semaphore_t** out_semaphore) {
  B200_REQUIRE_ARG(device && out_semaphore);
  auto* s = new iree_b200_semaphore_t();
  s->device = device;
  s->host_value = initial_value;
  iree_b200_device_retain(device);
  *out_semaphore = s;
  return IREE_B200_STATUS_OK;
}

void iree_b200_semaphore_release(iree_b200_semaphore_t* semaphore) {
  if (!semaphore || semaphore->ref_count.fetch_sub(1) != 1) return;
  iree_b200_device_t* device = semaphore->device;
  device->cu->cuCtxSetCurrent(device->cu_context);
  for (auto& tp : semaphore->timepoints) device->cu->cuEventDestroy(tp.event);
  delete semaphore;
  iree_b200_device_release(device);
}

// Retires timepoints whose events have completed. Caller holds semaphore->mutex.
static void semaphore_poll(iree_b200_semaphore_t* s) {
  const CudaSymbols* cu = s->device->cu;
  size_t retired = 0;
  for (auto& tp : s->timepoints) {
    CUresult r = cu->cuEventQuery(tp.event);
    if (r == CUDA_SUCCESS) {
      if (tp.value > s->host_value) s->host_value = tp.value;
      cu->cuEventDestroy(tp.event);
      ++retired;
    } else if (r == CUDA_ERROR_NOT_READY) {
      break;
    } else {
      if (s->failure == IREE_B200_STATUS_OK) {
        s->failure = cu_status(cu, r, "cuEventQuery", __FILE__, __LINE__);
        s->failure_message = iree_b200_status_message();
      }
      break;
    }
  }
  s->timepoints.erase(s->timepoints.begin(), s->timepoints.begin() + retired);
}

iree_b200_status_t iree_b200_semaphore_query(iree_b200_semaphore_t* semaphore, uint64_t* out_value) {
  B200_REQUIRE_ARG(semaphore && out_value);
  std::lock_guard<std::mutex> lock(semaphore->mutex);
  semaphore->device->cu->cuCtxSetCurrent(semaphore->device->cu_context);
  semaphore_poll(semaphore);
  *out_value = semaphore->host_value;
  if (semaphore->failure != IREE_B200_STATUS_OK) {
    return make_status(semaphore->failure, "%s", semaphore->failure_message.c_str());
  }
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t iree_b200_semaphore_signal(iree_b200_semaphore_t* semaphore, uint64_t value) {
  B200_REQUIRE_ARG(semaphore);
  std::lock_guard<std::mutex> lock(semaphore->mutex);
  semaphore->device->cu->cuCtxSetCurrent(semaphore->device->cu_context);
  semaphore_poll(semaphore);
  if (semaphore->failure != IREE_B200_STATUS_OK) {
    return make_status(semaphore->failure, "%s", semaphore->failure_message.c_str());
  }
  // Each timeline value is signalled once and values only grow (async/semaphore.c:577-599).
  if (value <= semaphore->host_value) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                       "semaphore already signaled past %llu (timeline is at %llu)",
                       (unsigned long long)value, (unsigned long long)semaphore->host_value);
  }
  semaphore->host_value = value;
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t iree_b200_semaphore_fail(iree_b200_semaphore_t* semaphore,
                                            iree_b200_status_t code, const char* message) {
  B200_REQUIRE_ARG(semaphore);
  if (code == IREE_B200_STATUS_OK) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "a semaphore cannot fail with OK");
  }
  std::lock_guard<std::mutex> lock(semaphore->mutex);
  if (semaphore->failure == IREE_B200_STATUS_OK) {  // the first failure wins
    semaphore->failure = code;
    semaphore->failure_message = message ? message : "semaphore failed by the host";
  }
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t iree_b200_semaphore_wait(iree_b200_semaphore_t* semaphore, uint64_t value,
                                            int64_t timeout_ns) {
  B200_REQUIRE_ARG(semaphore);
  const CudaSymbols* cu = semaphore->device->cu;
  const auto deadline = std::chrono::steady_clock::now() + std::chrono::nanoseconds(timeout_ns);
  for (;;) {
    CUevent event = nullptr;
    bool reached = false;
    {
      std::lock_guard<std::mutex> lock(semaphore->mutex);
      cu->cuCtxSetCurrent(semaphore->device->cu_context);
      semaphore_poll(semaphore);
      if (semaphore->failure != IREE_B200_STATUS_OK) {
        return make_status(semaphore->failure, "%s", semaphore->failure_message.c_str());
      }
      reached = semaphore->host_value >= value;
      for (auto& tp : semaphore->timepoints) {
        if (reached) break;
        if (tp.value >= value) {
          event = tp.event;
          break;
        }
      }
      if (event && timeout_ns < 0) {
        // Events are destroyed only under the mutex by poll(); synchronize while holding
        // it so the event cannot be retired underneath us.
        CUresult r = cu->cuEventSynchronize(event);
        if (r != CUDA_SUCCESS) {
          semaphore->failure = cu_status(cu, r, "cuEventSynchronize", __FILE__, __LINE__);
          semaphore->failure_message = iree_b200_status_message();
          return semaphore->failure;
        }
        continue;
      }
    }
    if (reached) {
      // Outside the semaphore's mutex (queue_execute takes device then semaphore): command
      // buffers whose submissions have completed are released here too.
      b200hal::reap_in_flight(semaphore->device, /*everything=*/false);
      return IREE_B200_STATUS_OK;
    }
    if (!event && timeout_ns < 0) {
      return make_status(IREE_B200_STATUS_FAILED_PRECONDITION,
                         "semaphore wait for %llu would never complete: no pending signal reaches it",
                         (unsigned long long)value);
    }
    if (timeout_ns >= 0 && std::chrono::steady_clock::now() >= deadline) {
      return make_status(IREE_B200_STATUS_DEADLINE_EXCEEDED, "semaphore wait timed out");
    }
    std::this_thread::sleep_for(std::chrono::microseconds(50));
  }
}

// Transient allocations in queue order. Like the reference without memory pools
// (cuda_device.c:1034-1091): the host waits for the wait list, allocates from the ordinary
// allocator, then host-signals the signal list — or fails it, so downstream waiters see the
// error instead of hanging.
static iree_b200_status_t wait_all(size_t count, iree_b200_semaphore_t* const* semaphores,
                                   const uint64_t* values) {
  for (size_t i = 0; i < count; ++i) {
    B200_RETURN_IF_ERROR(iree_b200_semaphore_wait(semaphores[i], values[i], /*timeout_ns=*/-1));
  }
  return IREE_B200_STATUS_OK;
}

static iree_b200_status_t signal_all(size_t count, iree_b200_semaphore_t* const* semaphores,
                                     const uint64_t* values) {
  for (size_t i = 0; i < count; ++i) {
    B200_RETURN_IF_ERROR(iree_b200_semaphore_signal(semaphores[i], values[i]));
  }
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t iree_b200_device_queue_alloca(
    iree_b200_device_t* device, size_t wait_count, iree_b200_semaphore_t* const* wait_semaphores,
    const uint64_t* wait_values, size_t signal_count,
    iree_b200_semaphore_t* const* signal_semaphores, const uint64_t* signal_values,
    int32_t memory_type, uint64_t byte_length, iree_b200_buffer_t** out_buffer) {
  B200_REQUIRE_ARG(device && out_buffer);
  B200_REQUIRE_ARG(!wait_count || (wait_semaphores && wait_values));
  B200_REQUIRE_ARG(!signal_count || (signal_semaphores && signal_values));
  *out_buffer = nullptr;
  B200_RETURN_IF_ERROR(wait_all(wait_count, wait_semaphores, wait_values));
  iree_b200_status_t status =
      iree_b200_allocator_allocate_buffer(device, memory_type, byte_length, out_buffer);
  if (status != IREE_B200_STATUS_OK) {
    const std::string message = iree_b200_status_message();
    for (size_t i = 0; i < signal_count; ++i) {
      iree_b200_semaphore_fail(signal_semaphores[i], status, message.c_str());
    }
    return make_status(status, "%s", message.c_str());
  }
  status = signal_all(signal_count, signal_semaphores, signal_values);
  if (status != IREE_B200_STATUS_OK) {
    const std::string message = iree_b200_status_message();
    iree_b200_buffer_release(*out_buffer);
    *out_buffer = nullptr;
    return make_status(status, "%s", message.c_str());
  }
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t iree_b200_device_queue_dealloca(
    iree_b200_device_t* device, size_t wait_count, iree_b200_semaphore_t* const* wait_semaphores,
    const uint64_t* wait_values, size_t signal_count,
    iree_b200_semaphore_t* const* signal_semaphores, const uint64_t* signal_values,
    iree_b200_buffer_t* buffer) {
  B200_REQUIRE_ARG(device && buffer);
  B200_REQUIRE_ARG(!wait_count || (wait_semaphores && wait_values));
  B200_REQUIRE_ARG(!signal_count || (signal_semaphores && signal_values));
  // The memory itself goes back when the last reference is released (cuda_device.c:1093-1140:
  // without pools dealloca is only a point on the timeline); the caller keeps its reference.
  B200_RETURN_IF_ERROR(wait_all(wait_count, wait_semaphores, wait_values));
  return signal_all(signal_count, signal_semaphores, signal_values);
}

}  // extern "C"
