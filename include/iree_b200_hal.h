// iree_b200_hal.h — C ABI of the thin B200 launch layer.
//
// This is the drop-in boundary for the run-time half of the dense-contraction dispatch
// path: the part of IREE's CUDA HAL driver that turns an executable blob plus
// (bindings, push constants, workgroup count) into kernel launches on a stream, and
// collective ops into NCCL calls. Every entry point names the reference interface it
// stands in for (paths relative to the iree-org/iree tree, runtime/src/iree/hal/...).
//
// Conventions (mirroring the reference):
//   * every fallible call returns iree_b200_status_t; 0 is OK, other values are
//     iree_status_code_t numbers (base/status.h:80-161). The message of the last
//     failure on the calling thread is available from iree_b200_status_message().
//     Nothing aborts on bad input; executable blobs are treated as untrusted
//     (drivers/cuda/native_executable.c:127-133).
//   * objects are reference counted; out_* parameters receive a +1 reference that the
//     caller drops with the matching *_release (hal/resource.h).
//   * plain C types only: pointers, sizes, fixed-width integers.
//   * libcuda / libnccl are dlopen'd on first use (drivers/cuda/cuda_dynamic_symbols.c:17-19,
//     nccl_dynamic_symbols.c), so this library loads on machines without a GPU; calls
//     that need the driver then fail with IREE_B200_STATUS_UNAVAILABLE. There is no CPU
//     fallback for any compute entry point.

#ifndef IREE_B200_HAL_H_
#define IREE_B200_HAL_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define IREE_B200_API
#else
#define IREE_B200_API __attribute__((visibility("default")))
#endif

//===----------------------------------------------------------------------===//
// Status (base/status.h:80-161)
//===----------------------------------------------------------------------===//

typedef int32_t iree_b200_status_t;
enum {
  IREE_B200_STATUS_OK = 0,
  IREE_B200_STATUS_CANCELLED = 1,
  IREE_B200_STATUS_UNKNOWN = 2,
  IREE_B200_STATUS_INVALID_ARGUMENT = 3,
  IREE_B200_STATUS_DEADLINE_EXCEEDED = 4,
  IREE_B200_STATUS_NOT_FOUND = 5,
  IREE_B200_STATUS_ALREADY_EXISTS = 6,
  IREE_B200_STATUS_PERMISSION_DENIED = 7,
  IREE_B200_STATUS_RESOURCE_EXHAUSTED = 8,
  IREE_B200_STATUS_FAILED_PRECONDITION = 9,
  IREE_B200_STATUS_ABORTED = 10,
  IREE_B200_STATUS_OUT_OF_RANGE = 11,
  IREE_B200_STATUS_UNIMPLEMENTED = 12,
  IREE_B200_STATUS_INTERNAL = 13,
  IREE_B200_STATUS_UNAVAILABLE = 14,
  IREE_B200_STATUS_DATA_LOSS = 15,
  IREE_B200_STATUS_INCOMPATIBLE = 18,
};

// Message of the most recent non-OK status returned on this thread ("" if none).
IREE_B200_API const char* iree_b200_status_message(void);
// "OK", "INVALID_ARGUMENT", ... (base/status.c iree_status_code_string).
IREE_B200_API const char* iree_b200_status_code_string(iree_b200_status_t status);

//===----------------------------------------------------------------------===//
// Limits (drivers/cuda/native_executable.h:24,28)
//===----------------------------------------------------------------------===//

#define IREE_B200_MAX_DISPATCH_BINDING_COUNT 16
#define IREE_B200_MAX_DISPATCH_CONSTANT_COUNT 64

typedef struct iree_b200_device_t iree_b200_device_t;
typedef struct iree_b200_buffer_t iree_b200_buffer_t;
typedef struct iree_b200_executable_cache_t iree_b200_executable_cache_t;
typedef struct iree_b200_executable_t iree_b200_executable_t;
typedef struct iree_b200_command_buffer_t iree_b200_command_buffer_t;
typedef struct iree_b200_semaphore_t iree_b200_semaphore_t;
typedef struct iree_b200_channel_t iree_b200_channel_t;

//===----------------------------------------------------------------------===//
// Device (drivers/cuda/api.h:59-124, cuda_device.c:411-420,555-639,772-804)
//===----------------------------------------------------------------------===//

enum {
  // cuda_device.c:417: the reference's programmatic default is graph mode; its tools
  // default to stream mode (registration/driver_module.c:16-19).
  IREE_B200_COMMAND_BUFFER_MODE_GRAPH = 0,
  IREE_B200_COMMAND_BUFFER_MODE_STREAM = 1,
};

typedef struct iree_b200_device_params_t {
  int32_t command_buffer_mode;  // IREE_B200_COMMAND_BUFFER_MODE_*
  int32_t reserved0;
  // When non-zero: an existing CUstream to use as the dispatch stream instead of
  // creating one (cuda_device.c:586). The caller keeps ownership.
  uint64_t external_dispatch_stream;
  // Collectives are issued on a second, high-priority stream and joined with events so
  // they can overlap independent dispatches (the reference issues NCCL on the single
  // dispatch stream, nccl_channel.c:542-588). 0 = reference behaviour. 1 = overlap; in
  // addition, symmetric-region all-reduces of equally sized, equally spaced ranges that
  // alternate with their producer dispatches between two barriers (GEMM chunk, all-reduce,
  // GEMM chunk, all-reduce, ...) are issued as ONE persistent launch: the kernel starts
  // behind the first producer and each later range is released by a stream write behind
  // its producer, so the ranks meet once per range instead of twice per launch and there
  // are no launch gaps. 2 = overlap without that fusion (diagnostics).
  // 3 = DATAFLOW: as 2, but an execution barrier no longer joins the two streams wholesale:
  // every command waits only for the latest command on the other stream, recorded before
  // the last barrier preceding it, with which it has a read/write hazard on its binding
  // ranges (read-only bindings, cuda_executable_def.fbs BindingBits, count as reads).
  // Kernels reach device memory through their bindings only, so the result is that of the
  // serialised execution; collectives then run under every dispatch that does not consume
  // them. The streams join at the end of the command buffer.
  int32_t overlap_collectives;
  int32_t reserved1;
} iree_b200_device_params_t;

// iree_hal_cuda_device_params_initialize (cuda_device.c:411-420).
IREE_B200_API void iree_b200_device_params_initialize(iree_b200_device_params_t* out_params);

// iree_hal_cuda_driver_create + create_device_by_id collapsed (api.h:120-124,
// cuda_device.c:555-639): retains the primary context of CUDA device |ordinal| and
// creates the dispatch stream.
IREE_B200_API iree_b200_status_t iree_b200_device_create(int32_t ordinal,
                                                         const iree_b200_device_params_t* params,
                                                         iree_b200_device_t** out_device);
IREE_B200_API void iree_b200_device_retain(iree_b200_device_t* device);
IREE_B200_API void iree_b200_device_release(iree_b200_device_t* device);

// iree_hal_cuda_device_query_i64 (cuda_device.c:772-804). Supported:
//   ("hal.device.id", "cuda") ("hal.executable.format", "cuda-nvptx-fb")
//   ("cuda.device", "compute_capability_major" | "compute_capability_minor")
//   ("b200.device", "multiprocessor_count" | "max_shared_memory_per_block_optin")
IREE_B200_API iree_b200_status_t iree_b200_device_query_i64(iree_b200_device_t* device,
                                                            const char* category, const char* key,
                                                            int64_t* out_value);

// The CUstream dispatches are issued on (as uint64).
IREE_B200_API uint64_t iree_b200_device_dispatch_stream(iree_b200_device_t* device);

// Blocks until everything queued on the device has finished (iree_hal_device_wait_idle).
IREE_B200_API iree_b200_status_t iree_b200_device_wait_idle(iree_b200_device_t* device);

//===----------------------------------------------------------------------===//
// Buffers (drivers/cuda/cuda_allocator.c:406-432,668; cuda_buffer.h:53)
//===----------------------------------------------------------------------===//

enum {
  IREE_B200_MEMORY_TYPE_DEVICE_LOCAL = 1,  // cuMemAlloc
  IREE_B200_MEMORY_TYPE_HOST_LOCAL = 2,    // cuMemHostAlloc (pinned, device visible)
};

IREE_B200_API iree_b200_status_t iree_b200_allocator_allocate_buffer(iree_b200_device_t* device,
                                                                     int32_t memory_type,
                                                                     uint64_t byte_length,
                                                                     iree_b200_buffer_t** out_buffer);
// Wraps device memory owned by someone else (e.g. a framework tensor); never freed here.
// Reference: iree_hal_cuda_allocator_import_buffer (cuda_allocator.c:668).
IREE_B200_API iree_b200_status_t iree_b200_allocator_import_buffer(iree_b200_device_t* device,
                                                                   uint64_t device_pointer,
                                                                   uint64_t byte_length,
                                                                   iree_b200_buffer_t** out_buffer);
IREE_B200_API void iree_b200_buffer_retain(iree_b200_buffer_t* buffer);
IREE_B200_API void iree_b200_buffer_release(iree_b200_buffer_t* buffer);
// iree_hal_cuda_buffer_device_pointer (cuda_buffer.h:53).
IREE_B200_API uint64_t iree_b200_buffer_device_pointer(const iree_b200_buffer_t* buffer);
IREE_B200_API void* iree_b200_buffer_host_pointer(const iree_b200_buffer_t* buffer);
IREE_B200_API uint64_t iree_b200_buffer_byte_length(const iree_b200_buffer_t* buffer);

// iree_hal_device_transfer_h2d / d2h (hal/device.h): asynchronous on the dispatch
// stream; |host| must stay valid until the device is idle or a semaphore signalled
// after the call has been waited on.
IREE_B200_API iree_b200_status_t iree_b200_device_transfer_h2d(iree_b200_device_t* device,
                                                               const void* host,
                                                               iree_b200_buffer_t* target,
                                                               uint64_t target_offset,
                                                               uint64_t byte_length);
IREE_B200_API iree_b200_status_t iree_b200_device_transfer_d2h(iree_b200_device_t* device,
                                                               iree_b200_buffer_t* source,
                                                               uint64_t source_offset, void* host,
                                                               uint64_t byte_length);

// A sub-range of a buffer (iree_hal_buffer_ref_t, hal/buffer.h).
typedef struct iree_b200_buffer_ref_t {
  iree_b200_buffer_t* buffer;
  uint64_t offset;
  uint64_t length;  // UINT64_MAX = to the end of the buffer
} iree_b200_buffer_ref_t;

//===----------------------------------------------------------------------===//
// Executables (drivers/cuda/nop_executable_cache.c:45-127, native_executable.c:95-477,
// schemas/cuda_executable_def.fbs, base/internal/flatcc/parsing.h:41-53)
//===----------------------------------------------------------------------===//

// iree_hal_cuda_nop_executable_cache_create (nop_executable_cache.c:45).
IREE_B200_API iree_b200_status_t iree_b200_executable_cache_create(
    iree_b200_device_t* device, const char* identifier, iree_b200_executable_cache_t** out_cache);
IREE_B200_API void iree_b200_executable_cache_release(iree_b200_executable_cache_t* cache);

// True for "PTXE" (nop_executable_cache.c:99-105).
IREE_B200_API int32_t iree_b200_executable_cache_can_prepare_format(
    iree_b200_executable_cache_t* cache, const char* executable_format);

// iree_hal_cuda_native_executable_infer_format (native_executable.c:95-125): checks the
// 64-byte header and the flatbuffer structure without touching the GPU; writes "PTXE".
// |data_length| 0 = unknown (size is then taken from the header).
IREE_B200_API iree_b200_status_t iree_b200_executable_infer_format(
    const void* executable_data, size_t data_length, size_t format_capacity,
    char* out_executable_format, size_t* out_inferred_size);

// iree_hal_cuda_native_executable_create (native_executable.c:250-477): verifies the
// "CDA1" flatbuffer against device limits, cuModuleLoadDataEx for every module,
// cuModuleGetFunction + MAX_DYNAMIC_SHARED_SIZE_BYTES for every export. In addition,
// looks up `<kernel_name>__tma` (include/iree_b200_tma_recipe.h) for TMA-fed kernels.
// The data is copied as needed; the caller may free it after the call returns.
IREE_B200_API iree_b200_status_t iree_b200_executable_cache_prepare_executable(
    iree_b200_executable_cache_t* cache, const char* executable_format,
    const void* executable_data, size_t data_length, iree_b200_executable_t** out_executable);
IREE_B200_API void iree_b200_executable_retain(iree_b200_executable_t* executable);
IREE_B200_API void iree_b200_executable_release(iree_b200_executable_t* executable);

// iree_hal_cuda_kernel_params_t (native_executable.h:36-55) as seen by callers.
typedef struct iree_b200_export_info_t {
  const char* kernel_name;  // owned by the executable
  uint32_t block_dims[3];
  uint32_t block_shared_memory_size;
  uint32_t constant_count;
  uint32_t binding_count;
  uint32_t tma_map_count;  // trailing CUtensorMap parameters (0 for plain kernels)
  uint32_t reserved;
} iree_b200_export_info_t;

IREE_B200_API int32_t iree_b200_executable_export_count(const iree_b200_executable_t* executable);
IREE_B200_API iree_b200_status_t iree_b200_executable_export_info(
    const iree_b200_executable_t* executable, int32_t export_ordinal,
    iree_b200_export_info_t* out_info);
// native_executable.c lookup_function_by_name.
IREE_B200_API iree_b200_status_t iree_b200_executable_lookup_export_by_name(
    const iree_b200_executable_t* executable, const char* name, int32_t* out_export_ordinal);

//===----------------------------------------------------------------------===//
// Command buffers (hal/command_buffer.h:1107-1178; drivers/cuda/stream_command_buffer.c,
// graph_command_buffer.c)
//===----------------------------------------------------------------------===//

// iree_hal_cuda_device_create_command_buffer (cuda_device.c:942-978). The device's
// command_buffer_mode picks the implementation: STREAM records a command list that is
// replayed onto the dispatch stream by queue_execute (stream_command_buffer.c:481-609
// behind hal/utils/deferred_command_buffer.c); GRAPH builds a CUgraph that is
// instantiated by end() and launched by queue_execute (graph_command_buffer.c:329-384).
IREE_B200_API iree_b200_status_t iree_b200_command_buffer_create(
    iree_b200_device_t* device, iree_b200_command_buffer_t** out_command_buffer);
IREE_B200_API void iree_b200_command_buffer_retain(iree_b200_command_buffer_t* command_buffer);
IREE_B200_API void iree_b200_command_buffer_release(iree_b200_command_buffer_t* command_buffer);

IREE_B200_API iree_b200_status_t iree_b200_command_buffer_begin(iree_b200_command_buffer_t* cb);
IREE_B200_API iree_b200_status_t iree_b200_command_buffer_end(iree_b200_command_buffer_t* cb);

// iree_hal_command_buffer_execution_barrier: commands recorded before the barrier
// complete before commands recorded after it start. Stream mode: stream order (no-op,
// stream_command_buffer.c:259-292) plus a flush of batched collectives (:154-171);
// graph mode: a dependency edge set (graph_command_buffer.c:413-439).
IREE_B200_API iree_b200_status_t iree_b200_command_buffer_execution_barrier(
    iree_b200_command_buffer_t* cb);

// iree_hal_command_buffer_fill_buffer (stream_command_buffer.c:327-367):
// |pattern_length| in {1,2,4}.
IREE_B200_API iree_b200_status_t iree_b200_command_buffer_fill_buffer(
    iree_b200_command_buffer_t* cb, iree_b200_buffer_ref_t target, const void* pattern,
    size_t pattern_length);
// iree_hal_command_buffer_update_buffer (:369-412): the source bytes are captured.
IREE_B200_API iree_b200_status_t iree_b200_command_buffer_update_buffer(
    iree_b200_command_buffer_t* cb, const void* source, size_t source_offset,
    iree_b200_buffer_ref_t target);
// iree_hal_command_buffer_copy_buffer (:414-453).
IREE_B200_API iree_b200_status_t iree_b200_command_buffer_copy_buffer(
    iree_b200_command_buffer_t* cb, iree_b200_buffer_ref_t source, iree_b200_buffer_ref_t target);

// iree_hal_command_buffer_dispatch (command_buffer.h:990-995; stream_command_buffer.c:481-609,
// graph_command_buffer.c:717-850). Kernel parameters are packed as the reference does:
// one device pointer per binding (buffer base + ref.offset), then each 32-bit constant
// in its own slot; for TMA-fed exports the encoded CUtensorMaps follow. Grid =
// |workgroup_count|, block = export block_dims, dynamic shared memory = export
// block_shared_memory_size. |constants| and |bindings| need not outlive the call.
IREE_B200_API iree_b200_status_t iree_b200_command_buffer_dispatch(
    iree_b200_command_buffer_t* cb, iree_b200_executable_t* executable, int32_t export_ordinal,
    const uint32_t workgroup_count[3], const uint32_t* constants, size_t constant_count,
    const iree_b200_buffer_ref_t* bindings, size_t binding_count, uint64_t flags);

// iree_hal_collective_op_t (command_buffer.h:300-462), same numeric encoding.
enum {
  IREE_B200_COLLECTIVE_KIND_ALL_GATHER = 0,
  IREE_B200_COLLECTIVE_KIND_ALL_REDUCE = 1,
  IREE_B200_COLLECTIVE_KIND_ALL_TO_ALL = 2,
  IREE_B200_COLLECTIVE_KIND_BROADCAST = 3,
  IREE_B200_COLLECTIVE_KIND_REDUCE = 4,
  IREE_B200_COLLECTIVE_KIND_REDUCE_SCATTER = 5,
  IREE_B200_COLLECTIVE_KIND_SEND = 6,
  IREE_B200_COLLECTIVE_KIND_RECV = 7,
  IREE_B200_COLLECTIVE_KIND_SEND_RECV = 8,
};
enum {
  IREE_B200_COLLECTIVE_REDUCTION_NONE = 0,
  IREE_B200_COLLECTIVE_REDUCTION_SUM = 1,
  IREE_B200_COLLECTIVE_REDUCTION_PRODUCT = 2,
  IREE_B200_COLLECTIVE_REDUCTION_MINIMUM = 3,
  IREE_B200_COLLECTIVE_REDUCTION_MAXIMUM = 4,
  IREE_B200_COLLECTIVE_REDUCTION_AVERAGE = 5,
};
enum {
  IREE_B200_COLLECTIVE_ELEMENT_TYPE_SINT_8 = 0,
  IREE_B200_COLLECTIVE_ELEMENT_TYPE_UINT_8 = 1,
  IREE_B200_COLLECTIVE_ELEMENT_TYPE_SINT_16 = 2,
  IREE_B200_COLLECTIVE_ELEMENT_TYPE_UINT_16 = 3,
  IREE_B200_COLLECTIVE_ELEMENT_TYPE_SINT_32 = 4,
  IREE_B200_COLLECTIVE_ELEMENT_TYPE_UINT_32 = 5,
  IREE_B200_COLLECTIVE_ELEMENT_TYPE_SINT_64 = 6,
  IREE_B200_COLLECTIVE_ELEMENT_TYPE_UINT_64 = 7,
  IREE_B200_COLLECTIVE_ELEMENT_TYPE_FLOAT_16 = 8,
  IREE_B200_COLLECTIVE_ELEMENT_TYPE_FLOAT_32 = 9,
  IREE_B200_COLLECTIVE_ELEMENT_TYPE_FLOAT_64 = 10,
  IREE_B200_COLLECTIVE_ELEMENT_TYPE_BFLOAT_16 = 11,
};
// packed = kind | reduction << 8 | element_type << 16 (little-endian struct layout of
// iree_hal_collective_op_t, command_buffer.h:446-462).
#define IREE_B200_COLLECTIVE_OP(kind, reduction, element_type) \
  ((uint32_t)(kind) | ((uint32_t)(reduction) << 8) | ((uint32_t)(element_type) << 16))

// iree_hal_command_buffer_collective (stream_command_buffer.c:455-479 → collective_batch
// → nccl_channel.c:327-588). Collectives recorded between two barriers are issued as one
// ncclGroupStart/End batch. BFLOAT_16 maps to ncclBfloat16 (the reference maps it to
// ncclFloat64, nccl_channel.c:291-293 — not reproduced).
// GRAPH mode: each batch is stream-captured at end() and embedded as a child-graph node
// ordered by the surrounding barriers (the reference returns UNIMPLEMENTED there and
// sketches this recipe in its TODO, graph_command_buffer.c:291-327). The symmetric-region
// kernels are a STREAM-mode path; in graphs those all-reduces go through NCCL.
IREE_B200_API iree_b200_status_t iree_b200_command_buffer_collective(
    iree_b200_command_buffer_t* cb, iree_b200_channel_t* channel, uint32_t packed_op,
    uint32_t param, iree_b200_buffer_ref_t send_binding, iree_b200_buffer_ref_t recv_binding,
    uint64_t element_count);

//===----------------------------------------------------------------------===//
// Semaphores and queue submission (cuda_device.c:1168-1196,260-281; event_semaphore.c)
//===----------------------------------------------------------------------===//

// Timeline semaphore backed by CUevents recorded on the dispatch stream.
IREE_B200_API iree_b200_status_t iree_b200_semaphore_create(iree_b200_device_t* device,
                                                            uint64_t initial_value,
                                                            iree_b200_semaphore_t** out_semaphore);
IREE_B200_API void iree_b200_semaphore_release(iree_b200_semaphore_t* semaphore);
IREE_B200_API iree_b200_status_t iree_b200_semaphore_query(iree_b200_semaphore_t* semaphore,
                                                           uint64_t* out_value);
// Host-side signal (iree_hal_semaphore_signal -> event_semaphore.c:108-131): advances the
// timeline to |value|. Values only grow: INVALID_ARGUMENT if the timeline is already at or
// past |value| (async/semaphore.c:577-599). Submissions and waits issued AFTER the signal see
// it; this thin layer has no deferred work queue, so a submission that waits on a value
// nothing queued or signalled reaches is refused rather than parked (see queue_execute).
IREE_B200_API iree_b200_status_t iree_b200_semaphore_signal(iree_b200_semaphore_t* semaphore,
                                                            uint64_t value);
// Marks the semaphore failed (iree_hal_semaphore_fail; async/semaphore.c:185-204): the first
// failure wins, and every later query / wait / signal / submission that involves the
// semaphore returns |code| with |message| (NULL = a default text). |code| must not be OK.
IREE_B200_API iree_b200_status_t iree_b200_semaphore_fail(iree_b200_semaphore_t* semaphore,
                                                          iree_b200_status_t code,
                                                          const char* message);
// Blocks the host until the timeline reaches |value| (cuEventSynchronize,
// cuda_device.c:184-191). |timeout_ns| < 0 = infinite.
IREE_B200_API iree_b200_status_t iree_b200_semaphore_wait(iree_b200_semaphore_t* semaphore,
                                                          uint64_t value, int64_t timeout_ns);

// iree_hal_device_queue_execute (cuda_device.c:1168-1196): waits (device side) for every
// (wait_semaphores[i] >= wait_values[i]), runs |command_buffer| (may be NULL), then
// advances every signal semaphore. Async failures fail the signal semaphores
// (cuda_device.c:1187-1191): a later wait returns the error.
IREE_B200_API iree_b200_status_t iree_b200_device_queue_execute(
    iree_b200_device_t* device, size_t wait_count, iree_b200_semaphore_t* const* wait_semaphores,
    const uint64_t* wait_values, size_t signal_count,
    iree_b200_semaphore_t* const* signal_semaphores, const uint64_t* signal_values,
    iree_b200_command_buffer_t* command_buffer);

// iree_hal_device_queue_alloca / _dealloca (cuda_device.c:1034-1140), the form the reference
// takes on devices without memory pools: the HOST waits for the wait list, the buffer comes
// from the ordinary allocator, and the signal list is then signalled from the host — or
// failed with the allocation's error, so that downstream waiters see it instead of hanging.
// dealloca is a point on the timeline only: the memory goes back when the last reference to
// |buffer| is released (the caller keeps its own).
IREE_B200_API iree_b200_status_t iree_b200_device_queue_alloca(
    iree_b200_device_t* device, size_t wait_count, iree_b200_semaphore_t* const* wait_semaphores,
    const uint64_t* wait_values, size_t signal_count,
    iree_b200_semaphore_t* const* signal_semaphores, const uint64_t* signal_values,
    int32_t memory_type, uint64_t byte_length, iree_b200_buffer_t** out_buffer);
IREE_B200_API iree_b200_status_t iree_b200_device_queue_dealloca(
    iree_b200_device_t* device, size_t wait_count, iree_b200_semaphore_t* const* wait_semaphores,
    const uint64_t* wait_values, size_t signal_count,
    iree_b200_semaphore_t* const* signal_semaphores, const uint64_t* signal_values,
    iree_b200_buffer_t* buffer);

//===----------------------------------------------------------------------===//
// Channels (hal/channel.h:33-129, channel_provider.h:29-66; drivers/cuda/nccl_channel.c,
// cuda_device.c:845-927)
//===----------------------------------------------------------------------===//

#define IREE_B200_CHANNEL_ID_SIZE 128  // sizeof(ncclUniqueId)
#define IREE_B200_CHANNEL_RANK_DEFAULT (-1)
#define IREE_B200_CHANNEL_COUNT_DEFAULT (-1)
#define IREE_B200_CHANNEL_NO_COLOR (-1)

typedef struct iree_b200_channel_params_t {
  uint64_t flags;
  const uint8_t* id;  // ncclUniqueId bytes or NULL: obtained through the default provider
  size_t id_length;   // 0 or IREE_B200_CHANNEL_ID_SIZE
  int32_t rank;       // or IREE_B200_CHANNEL_RANK_DEFAULT
  int32_t count;      // or IREE_B200_CHANNEL_COUNT_DEFAULT
} iree_b200_channel_params_t;

// iree_hal_cuda_nccl_get_unique_id (nccl_channel.c:76-92).
IREE_B200_API iree_b200_status_t iree_b200_channel_id_create(
    uint8_t out_id[IREE_B200_CHANNEL_ID_SIZE]);

// Default channel provider (channel_provider.h:57-66) without MPI: rank/count from
// IREE_B200_RANK / IREE_B200_WORLD_SIZE, else RANK / WORLD_SIZE (torchrun), else
// OMPI_COMM_WORLD_RANK / _SIZE (mpi_channel_provider.c:164-173); the id is exchanged
// through files next to the path named by IREE_B200_CHANNEL_ID_FILE, standing in for
// MPI_Bcast (mpi_channel_provider.c:185-190): every rank publishes a fresh random token,
// rank 0 publishes {sequence number, tokens, id}, a rank accepts only the id file that
// carries its own token and acknowledges it, rank 0 returns when all have and removes the
// files — so a file left by an earlier run or an earlier channel is never taken for the id.
// Like MPI_Bcast it is collective: all ranks call it, in the same order per channel.
IREE_B200_API iree_b200_status_t iree_b200_channel_provider_query_default_rank_and_count(
    int32_t* out_rank, int32_t* out_count);
IREE_B200_API iree_b200_status_t iree_b200_channel_provider_exchange_default_id(
    int32_t rank, uint8_t inout_id[IREE_B200_CHANNEL_ID_SIZE]);

// iree_hal_cuda_device_create_channel (cuda_device.c:845-927) →
// iree_hal_cuda_nccl_channel_create (nccl_channel.c:94-140, ncclCommInitRankConfig).
IREE_B200_API iree_b200_status_t iree_b200_channel_create(iree_b200_device_t* device,
                                                          const iree_b200_channel_params_t* params,
                                                          iree_b200_channel_t** out_channel);
IREE_B200_API void iree_b200_channel_retain(iree_b200_channel_t* channel);
IREE_B200_API void iree_b200_channel_release(iree_b200_channel_t* channel);
// iree_hal_channel_split (nccl_channel.c:176-231, ncclCommSplit). *out is NULL when
// |color| is IREE_B200_CHANNEL_NO_COLOR.
IREE_B200_API iree_b200_status_t iree_b200_channel_split(iree_b200_channel_t* base_channel,
                                                         int32_t color, int32_t key,
                                                         uint64_t flags,
                                                         iree_b200_channel_t** out_split_channel);
IREE_B200_API void iree_b200_channel_query_rank_and_count(const iree_b200_channel_t* channel,
                                                          int32_t* out_rank, int32_t* out_count);

// Symmetric region: device memory every rank of the channel has mapped at once (CUDA VMM
// peer mappings + an NVSwitch multicast object; the host side sets those up, e.g. with
// torch.distributed._symmetric_memory — plumbing). Once attached, an in-place
// all_reduce(sum) of f16 / bf16 / f32 recorded with iree_b200_command_buffer_collective
// whose range lies inside the region (16-byte aligned, a multiple of 16 bytes) is executed
// by the builtin NVLink kernels (kernels/collective_kernels.cu: multimem.ld_reduce +
// multimem.st through the switch, or peer loads/stores when there is no multicast
// mapping) instead of ncclAllReduce; every other collective still goes to NCCL. All ranks
// must attach the same region and record the same collectives in the same order — the
// contract iree_hal_channel already has (hal/channel.h:33-60).
#define IREE_B200_SYMMETRIC_MAX_RANKS 8
#define IREE_B200_SYMMETRIC_FLAG_BYTES 2048  // per rank, zero-initialised, inside or outside the region

typedef struct iree_b200_symmetric_region_params_t {
  uint64_t byte_length;
  // This rank's mappings of every rank's copy of the region; peer_bases[rank] is the local one.
  const uint64_t* peer_bases;
  // Multicast mapping of the region, or 0: then the peer-load/store kernels are used.
  uint64_t multicast_base;
  // This rank's mappings of every rank's flag block (IREE_B200_SYMMETRIC_FLAG_BYTES each).
  const uint64_t* peer_flag_blocks;
  // Upper bound on the CTAs one collective uses (<= 64; 0 = default 16). Collectives overlap
  // with persistent GEMM grids only if that many SMs are left free
  // (iree_b200_ops_set_reserved_sms).
  int32_t max_blocks;
} iree_b200_symmetric_region_params_t;

// |executable| must export the iree_b200_all_reduce_sum_* kernels (the builtin one does:
// iree_b200_ops_executable).
IREE_B200_API iree_b200_status_t iree_b200_channel_attach_symmetric_region(
    iree_b200_channel_t* channel, iree_b200_executable_t* executable,
    const iree_b200_symmetric_region_params_t* params);
// Tuning and watchdog of the symmetric kernels are read from the environment at attach
// time (every rank must see the same values):
//   IREE_B200_ALL_REDUCE_TUNING="threads,unroll,pipelined"   block size (multiple of 32,
//       <= 1024), 16-byte multimem requests in flight per thread (1 / 2 / 4 / 8) and whether
//       the loads of step i+1 are issued before the stores of step i; default "1024,2,1".
//   IREE_B200_COLLECTIVE_WATCHDOG_MS=<n>   a wait on a peer (or on a producer in a fused
//       range launch) that lasts longer makes the launch record the reason and retire (no
//       trap: the context stays usable; the channel must then be considered failed), readable
//       with iree_b200_channel_query_collective_status. Default 0: wait indefinitely, as a
//       NCCL collective does (ranks may legitimately be seconds apart: first-launch module
//       load, a profiler replay, a long producer).
// Number of collectives this channel has issued through the symmetric kernels / NCCL.
IREE_B200_API void iree_b200_channel_query_collective_counts(const iree_b200_channel_t* channel,
                                                             uint64_t* out_symmetric,
                                                             uint64_t* out_nccl);

// 0 while no symmetric kernel of this channel has timed out; otherwise reason (1 = a put to
// a peer's flag word never succeeded, 2 = a peer never arrived, 3 = a fused range was never
// released by its producer) | block index << 8. Reads device memory (synchronous, tiny).
IREE_B200_API iree_b200_status_t iree_b200_channel_query_collective_status(
    iree_b200_channel_t* channel, uint32_t* out_status);

#ifdef __cplusplus
}  // extern "C"
#endif

#endif  // IREE_B200_HAL_H_
