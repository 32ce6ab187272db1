// internal.h — shared internals of the thin B200 launch layer (not part of the ABI).

#ifndef IREE_B200_HAL_INTERNAL_H_
#define IREE_B200_HAL_INTERNAL_H_

#include <cuda.h>
#include <nccl.h>

#include <atomic>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "iree_b200_hal.h"
#include "iree_b200_tma_recipe.h"

namespace b200hal {

//===----------------------------------------------------------------------===//
// Status
//===----------------------------------------------------------------------===//

// Records |fmt| as the calling thread's last message and returns |code|.
iree_b200_status_t make_status(iree_b200_status_t code, const char* fmt, ...)
    __attribute__((format(printf, 2, 3)));

#define B200_RETURN_IF_ERROR(expr)                        \
  do {                                                    \
    iree_b200_status_t b200_status__ = (expr);            \
    if (b200_status__ != IREE_B200_STATUS_OK) return b200_status__; \
  } while (0)

#define B200_REQUIRE_ARG(cond)                                                          \
  do {                                                                                  \
    if (!(cond))                                                                        \
      return ::b200hal::make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "%s: `%s` violated", \
                                    __func__, #cond);                                   \
  } while (0)

//===----------------------------------------------------------------------===//
// Dynamic symbols (drivers/cuda/cuda_dynamic_symbols.c, nccl_dynamic_symbols.c)
//===----------------------------------------------------------------------===//

// X-macro list; cuda.h maps the unsuffixed names to their _v2 forms, and the loader
// stringifies AFTER macro expansion so the versioned symbol is looked up.
#define B200_CUDA_SYMBOLS(X)       \
  X(cuInit)                        \
  X(cuDriverGetVersion)            \
  X(cuGetErrorName)                \
  X(cuGetErrorString)              \
  X(cuDeviceGet)                   \
  X(cuDeviceGetCount)              \
  X(cuDeviceGetName)               \
  X(cuDeviceGetAttribute)          \
  X(cuDevicePrimaryCtxRetain)      \
  X(cuDevicePrimaryCtxRelease)     \
  X(cuCtxSetCurrent)               \
  X(cuCtxGetCurrent)               \
  X(cuCtxSynchronize)              \
  X(cuStreamCreateWithPriority)    \
  X(cuStreamCreate)                \
  X(cuStreamDestroy)               \
  X(cuStreamSynchronize)           \
  X(cuStreamWaitEvent)             \
  X(cuStreamWriteValue32)          \
  X(cuCtxGetStreamPriorityRange)   \
  X(cuEventCreate)                 \
  X(cuEventDestroy)                \
  X(cuEventRecord)                 \
  X(cuEventQuery)                  \
  X(cuEventSynchronize)            \
  X(cuEventElapsedTime)            \
  X(cuMemAlloc)                    \
  X(cuMemFree)                     \
  X(cuMemHostAlloc)                \
  X(cuMemFreeHost)                 \
  X(cuMemHostGetDevicePointer)     \
  X(cuMemcpyHtoDAsync)             \
  X(cuMemcpyDtoHAsync)             \
  X(cuMemcpyAsync)                 \
  X(cuMemcpy2DAsync)               \
  X(cuMemsetD8Async)               \
  X(cuMemsetD16Async)              \
  X(cuMemsetD32Async)              \
  X(cuModuleLoadDataEx)            \
  X(cuModuleUnload)                \
  X(cuModuleGetFunction)           \
  X(cuModuleGetGlobal)             \
  X(cuFuncSetAttribute)            \
  X(cuFuncGetAttribute)            \
  X(cuLaunchKernel)                \
  X(cuTensorMapEncodeTiled)        \
  X(cuGraphCreate)                 \
  X(cuGraphDestroy)                \
  X(cuGraphAddKernelNode)          \
  X(cuGraphAddMemsetNode)          \
  X(cuGraphAddMemcpyNode)          \
  X(cuGraphAddEmptyNode)           \
  X(cuGraphAddChildGraphNode)      \
  X(cuStreamBeginCapture)          \
  X(cuStreamEndCapture)            \
  X(cuGraphInstantiateWithFlags)   \
  X(cuGraphExecDestroy)            \
  X(cuGraphLaunch)

#define B200_NCCL_SYMBOLS(X) \
  X(ncclGetVersion)          \
  X(ncclGetUniqueId)         \
  X(ncclCommInitRankConfig)  \
  X(ncclCommInitRank)        \
  X(ncclCommSplit)           \
  X(ncclCommDestroy)         \
  X(ncclCommAbort)           \
  X(ncclCommCount)           \
  X(ncclCommUserRank)        \
  X(ncclGetErrorString)      \
  X(ncclGetLastError)        \
  X(ncclAllGather)           \
  X(ncclAllReduce)           \
  X(ncclBroadcast)           \
  X(ncclReduce)              \
  X(ncclReduceScatter)       \
  X(ncclSend)                \
  X(ncclRecv)                \
  X(ncclGroupStart)          \
  X(ncclGroupEnd)

struct CudaSymbols {
#define B200_DECL(name) decltype(&::name) name = nullptr;
  B200_CUDA_SYMBOLS(B200_DECL)
#undef B200_DECL
};
struct NcclSymbols {
#define B200_DECL(name) decltype(&::name) name = nullptr;
  B200_NCCL_SYMBOLS(B200_DECL)
#undef B200_DECL
};

// Loads libcuda.so.1 once and calls cuInit(0). UNAVAILABLE when there is no driver.
iree_b200_status_t cuda_symbols(const CudaSymbols** out_syms);
// Loads libnccl.so.2 once.
iree_b200_status_t nccl_symbols(const NcclSymbols** out_syms);

// CUresult / ncclResult_t → status with the driver's error name (cuda_status_util.h:24-55).
iree_b200_status_t cu_status(const CudaSymbols* syms, CUresult result, const char* expr,
                             const char* file, int line);
iree_b200_status_t nccl_status(const NcclSymbols* syms, ncclResult_t result, const char* expr,
                               const char* file, int line);

#define B200_CU(syms, call)                                                             \
  do {                                                                                  \
    CUresult b200_cu__ = (syms)->call;                                                  \
    if (b200_cu__ != CUDA_SUCCESS)                                                      \
      return ::b200hal::cu_status((syms), b200_cu__, #call, __FILE__, __LINE__);        \
  } while (0)

#define B200_NCCL(syms, call)                                                           \
  do {                                                                                  \
    ncclResult_t b200_nccl__ = (syms)->call;                                            \
    if (b200_nccl__ != ncclSuccess)                                                     \
      return ::b200hal::nccl_status((syms), b200_nccl__, #call, __FILE__, __LINE__);    \
  } while (0)

//===----------------------------------------------------------------------===//
// Objects
//===----------------------------------------------------------------------===//

struct RefCounted {
  std::atomic<int32_t> ref_count{1};
};

}  // namespace b200hal

struct iree_b200_device_t : b200hal::RefCounted {
  const b200hal::CudaSymbols* cu = nullptr;
  iree_b200_device_params_t params{};
  CUdevice cu_device = 0;
  CUcontext cu_context = nullptr;
  CUstream dispatch_stream = nullptr;
  bool owns_dispatch_stream = false;
  CUstream collective_stream = nullptr;  // only with params.overlap_collectives
  int32_t ordinal = 0;
  int32_t cc_major = 0, cc_minor = 0;
  int32_t sm_count = 0;
  int32_t max_smem_optin = 0;
  int32_t max_block_dims[3] = {0, 0, 0};
  int32_t max_threads_per_block = 0;
  std::mutex mutex;  // guards stream submission order
  // Command buffers the GPU may still be executing: retained by queue_execute until the
  // event recorded behind them has fired (the reference keeps them until the signal
  // timepoint, cuda_device.c:1168-1196). Guarded by |mutex|; reaped by queue_execute,
  // wait_idle and host semaphore waits.
  struct InFlight {
    iree_b200_command_buffer_t* command_buffer;
    CUevent done;
  };
  std::vector<InFlight> in_flight;
};

namespace b200hal {
// Releases the in-flight command buffers whose work has completed (all of them when
// |everything|: the caller has just synchronised the streams). Takes the device mutex.
void reap_in_flight(iree_b200_device_t* device, bool everything);
}  // namespace b200hal

struct iree_b200_buffer_t : b200hal::RefCounted {
  iree_b200_device_t* device = nullptr;
  int32_t memory_type = 0;
  bool imported = false;
  CUdeviceptr device_ptr = 0;
  void* host_ptr = nullptr;
  uint64_t byte_length = 0;
};

struct iree_b200_executable_cache_t : b200hal::RefCounted {
  iree_b200_device_t* device = nullptr;
  std::string identifier;
};

namespace b200hal {

// iree_hal_cuda_kernel_params_t (native_executable.h:36-55) + the TMA recipe.
struct KernelParams {
  std::string name;
  CUfunction function = nullptr;
  uint32_t block_dims[3] = {1, 1, 1};
  uint32_t block_shared_memory_size = 0;
  uint32_t constant_count = 0;
  uint32_t binding_count = 0;
  uint64_t binding_flags[IREE_B200_MAX_DISPATCH_BINDING_COUNT] = {0};
  bool has_recipe = false;
  iree_b200_tma_recipe_t recipe{};
};

// Parsed view of a CDA1 blob (schemas/cuda_executable_def.fbs). Pure host logic.
struct ExportDefView {
  uint32_t module_ordinal = 0;
  std::string kernel_name;
  bool has_block_dims = false;
  uint32_t block_dims[3] = {0, 0, 0};
  uint32_t block_shared_memory_size = 0;
  uint32_t constant_count = 0;
  std::vector<uint64_t> binding_flags;
  bool present = true;
};
struct ExecutableDefView {
  std::vector<ExportDefView> exports;
  std::vector<std::pair<const uint8_t*, size_t>> modules;  // ptx_image bytes
};

// Checks the 64-byte header (hal/utils/executable_header.c:9-72).
iree_b200_status_t read_executable_header(const void* data, size_t data_length,
                                          bool unsafe_infer_size, const uint8_t** out_fb,
                                          size_t* out_fb_length);
// Bounds-checked flatbuffer walk (the job of the flatcc-generated verifier + reader).
iree_b200_status_t parse_executable_def(const uint8_t* fb, size_t fb_length,
                                        ExecutableDefView* out_view);

// Fills |out_map| (CUtensorMap, 128 bytes) from a recipe entry.
iree_b200_status_t encode_tensor_map(const CudaSymbols* cu, const iree_b200_tma_map_t& map,
                                     uint64_t base_pointer, const uint32_t* constants,
                                     size_t constant_count, CUtensorMap* out_map);

}  // namespace b200hal

struct iree_b200_executable_t : b200hal::RefCounted {
  iree_b200_device_t* device = nullptr;
  std::vector<CUmodule> modules;
  std::vector<b200hal::KernelParams> exports;
};

struct iree_b200_channel_t : b200hal::RefCounted {
  iree_b200_device_t* device = nullptr;
  const b200hal::NcclSymbols* nccl = nullptr;
  ncclComm_t comm = nullptr;
  int32_t rank = 0;
  int32_t count = 1;
  // Symmetric region (iree_b200_channel_attach_symmetric_region).
  struct {
    bool attached = false;
    CUdeviceptr local_base = 0;
    uint64_t byte_length = 0;
    CUdeviceptr multicast_base = 0;
    CUdeviceptr table = 0;  // device copy of {peer bases[8], peer flag blocks[8]}, then the
                            // "range ready" word of fused multi-range launches (offset 128)
    uint32_t ready_sequence = 0;  // last value written to / expected in that word
    int32_t max_blocks = 0;
    int32_t threads = 0;   // block size of the all-reduce launches
    uint32_t tuning = 0;   // kernel |tuning| parameter (unroll | pipelined << 8)
    iree_b200_executable_t* executable = nullptr;
    CUfunction all_reduce_sum[3] = {nullptr, nullptr, nullptr};  // f16, bf16, f32
  } symmetric;
  std::atomic<uint64_t> symmetric_collectives{0}, nccl_collectives{0};
};

struct iree_b200_semaphore_t : b200hal::RefCounted {
  iree_b200_device_t* device = nullptr;
  std::mutex mutex;
  uint64_t host_value = 0;  // highest value known reached without an event
  struct Timepoint {
    uint64_t value;
    CUevent event;
  };
  std::vector<Timepoint> timepoints;  // ascending by value
  iree_b200_status_t failure = IREE_B200_STATUS_OK;
  std::string failure_message;
};

#endif  // IREE_B200_HAL_INTERNAL_H_
