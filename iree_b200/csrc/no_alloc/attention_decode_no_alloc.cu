// attention_decode_no_alloc.cu — iree_b200_attention_decode_no_alloc: split-KV decode attention
// whose scratch memory is a static device array, for callers that cannot allocate.
//
// No counterpart in the reference. Some callers
// cannot call cudaMalloc — code under CUDA graph capture, tight serving loops, harnesses that
// forbid it — and the split-KV decode pair (attention_kernels.cu: iree_b200_attention_decode_
// split_* + _merge_*) needs a workspace of partial softmax records between its two launches,
// which every other entry point takes from a caller buffer or from the ops object's arena.
// Here the records live in |g_decode_scratch|, a file-scope `static __device__` array of
// IREE_B200_NO_ALLOC_SCRATCH_BYTES: every GPU that loads this library holds that much until
// the process exits. The entry point calls no cudaMalloc / cudaFree / allocating routine.
//
// Unlike the rest of the library (cubins in a CDA1 container, launched through the driver
// API), this one translation unit is an nvcc object with the CUDA runtime linked statically:
// a static __device__ symbol and cudaGetSymbolAddress belong to the runtime's module
// registration. The kernels are the SAME device bodies as the exported pair
// (kernels/attention_decode_split.cuh), so the results are bit-identical to it.
//
// Sharing: every call, on every stream, uses the one array of its device. Calls that overlap
// on the GPU race on it; callers serialise them (one stream, or events). The array keeps the
// last call's contents: nothing has to start at zero, because the split kernel writes every
// field of every record the merge kernel reads (empty splits write max = -FLT_MAX, sum = 0),
// so no clearing pass is needed — the test runs two differently-shaped calls back to back.

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <mutex>

#include "../kernels/attention_decode_split.cuh"
#include "iree_b200_ops.h"

namespace b200hal {
// hal/device.cc: records the calling thread's last message and returns |code|.
iree_b200_status_t make_status(iree_b200_status_t code, const char* fmt, ...)
    __attribute__((format(printf, 2, 3)));
}  // namespace b200hal

namespace {

using b200::AttnDecodeConfig;

constexpr size_t kScratchBytes = IREE_B200_NO_ALLOC_SCRATCH_BYTES;
constexpr size_t kRecordBytes = AttnDecodeConfig::kRecordFloats * sizeof(float);
constexpr int kMaxDevices = 64;

// The scratch. In HOST code this name is a host-side stand-in, not a device address: the
// address the kernels get comes from cudaGetSymbolAddress (scratch_of below).
static __device__ __align__(256) float g_decode_scratch[kScratchBytes / sizeof(float)];

template <bool kBf16>
__global__ void __launch_bounds__(AttnDecodeConfig::kThreads)
    decode_split_no_alloc(const uint16_t* __restrict__ q, const uint16_t* __restrict__ k,
                          const uint16_t* __restrict__ v, float* __restrict__ scratch, int B, int M,
                          int K2, int scale_bits, int splits) {
  b200::attention_decode_split_body<kBf16>(q, k, v, scratch, B, M, K2, scale_bits, splits);
}

template <bool kBf16>
__global__ void __launch_bounds__(128)
    decode_merge_no_alloc(const float* __restrict__ scratch, uint16_t* __restrict__ out, int B,
                          int M, int splits) {
  b200::attention_decode_merge_body<kBf16>(scratch, out, B, M, splits);
}

struct DeviceState {
  float* scratch = nullptr;  // device address of g_decode_scratch on this device
  int sm_count = 0;
};
std::mutex g_mutex;
DeviceState g_devices[kMaxDevices];

// Set-up, once per device, with that device current.
cudaError_t scratch_of(int device, DeviceState* out) {
  std::lock_guard<std::mutex> lock(g_mutex);
  DeviceState& state = g_devices[device];
  if (!state.scratch) {
    float* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_decode_scratch);
    if (e != cudaSuccess) return e;
    int sms = 0;
    e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    if (e != cudaSuccess) return e;
    state.sm_count = sms;
    state.scratch = ptr;
  }
  *out = state;
  return cudaSuccess;
}

}  // namespace

extern "C" IREE_B200_API iree_b200_status_t iree_b200_attention_decode_no_alloc(
    int32_t device_ordinal, int32_t element_type, int64_t B, int64_t M, int64_t K2, float scale,
    const void* q, const void* k, const void* v, void* out, void* stream) {
  using b200hal::make_status;
  if (element_type != IREE_B200_ELEMENT_TYPE_F16 && element_type != IREE_B200_ELEMENT_TYPE_BF16) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                       "attention_decode_no_alloc takes f16 or bf16 (got element type %d)",
                       element_type);
  }
  if (device_ordinal < 0 || device_ordinal >= kMaxDevices) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "device ordinal %d out of range",
                       device_ordinal);
  }
  if (B < 0 || M < 0 || K2 < 1 || B > 65535 || M > 65535 || K2 > INT32_MAX) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                       "attention_decode_no_alloc shape B=%lld M=%lld K2=%lld out of range "
                       "(B, M <= 65535; 1 <= K2 < 2^31)",
                       (long long)B, (long long)M, (long long)K2);
  }
  if (B == 0 || M == 0) return IREE_B200_STATUS_OK;  // empty result
  if ((uint64_t)B * M * kRecordBytes > kScratchBytes) {  // even one split per row does not fit
    return make_status(IREE_B200_STATUS_RESOURCE_EXHAUSTED,
                       "attention_decode_no_alloc: B=%lld M=%lld needs at least %llu bytes of scratch "
                       "but the static array holds IREE_B200_NO_ALLOC_SCRATCH_BYTES = %zu; use "
                       "iree_b200_ops_record_attention with a caller workspace instead",
                       (long long)B, (long long)M, (unsigned long long)((uint64_t)B * M * kRecordBytes),
                       kScratchBytes);
  }
  if (!q || !k || !v || !out) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "attention_decode_no_alloc: null operand");
  }
  if (((uintptr_t)q | (uintptr_t)k | (uintptr_t)v | (uintptr_t)out) % 16) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                       "attention_decode_no_alloc: operands must be 16-byte aligned");
  }

  int previous = -1;
  cudaError_t e = cudaGetDevice(&previous);
  if (e == cudaSuccess && previous != device_ordinal) e = cudaSetDevice(device_ordinal);
  DeviceState state;
  if (e == cudaSuccess) e = scratch_of(device_ordinal, &state);
  iree_b200_status_t status = IREE_B200_STATUS_OK;
  if (e != cudaSuccess) {
    status = make_status(IREE_B200_STATUS_UNAVAILABLE,
                         "attention_decode_no_alloc: set-up on device %d failed: %s", device_ordinal,
                         cudaGetErrorString(e));
  } else {
    // The split count of iree_b200_ops_select_attention's split + merge form (ops.cc): enough
    // CTAs for ~8 per SM, at most one per 256-row chunk of K2.
    const int64_t chunks = (K2 + AttnDecodeConfig::kChunkRows - 1) / AttnDecodeConfig::kChunkRows;
    const int64_t want = state.sm_count * 8LL / (B * M) > 1 ? state.sm_count * 8LL / (B * M) : 1;
    const int64_t splits = chunks < want ? chunks : want;
    const uint64_t need = (uint64_t)B * M * splits * kRecordBytes;
    if (need > kScratchBytes) {
      status = make_status(IREE_B200_STATUS_RESOURCE_EXHAUSTED,
                           "attention_decode_no_alloc: B=%lld M=%lld K2=%lld needs %llu bytes of "
                           "scratch (%lld splits x %zu bytes per row) but the static array holds "
                           "IREE_B200_NO_ALLOC_SCRATCH_BYTES = %zu; use "
                           "iree_b200_ops_record_attention with a caller workspace instead",
                           (long long)B, (long long)M, (long long)K2, (unsigned long long)need,
                           (long long)splits, kRecordBytes, kScratchBytes);
    } else {
      uint32_t scale_bits;
      memcpy(&scale_bits, &scale, 4);
      const dim3 split_grid((unsigned)splits, (unsigned)M, (unsigned)B);
      const dim3 merge_grid((unsigned)M, (unsigned)B, 1);
      cudaStream_t s = static_cast<cudaStream_t>(stream);
      const uint16_t* q16 = static_cast<const uint16_t*>(q);
      const uint16_t* k16 = static_cast<const uint16_t*>(k);
      const uint16_t* v16 = static_cast<const uint16_t*>(v);
      uint16_t* out16 = static_cast<uint16_t*>(out);
      if (element_type == IREE_B200_ELEMENT_TYPE_BF16) {
        decode_split_no_alloc<true><<<split_grid, AttnDecodeConfig::kThreads, 0, s>>>(
            q16, k16, v16, state.scratch, (int)B, (int)M, (int)K2, (int)scale_bits, (int)splits);
        decode_merge_no_alloc<true><<<merge_grid, 128, 0, s>>>(state.scratch, out16, (int)B, (int)M,
                                                               (int)splits);
      } else {
        decode_split_no_alloc<false><<<split_grid, AttnDecodeConfig::kThreads, 0, s>>>(
            q16, k16, v16, state.scratch, (int)B, (int)M, (int)K2, (int)scale_bits, (int)splits);
        decode_merge_no_alloc<false><<<merge_grid, 128, 0, s>>>(state.scratch, out16, (int)B, (int)M,
                                                                (int)splits);
      }
      e = cudaGetLastError();
      if (e != cudaSuccess) {
        status = make_status(IREE_B200_STATUS_INTERNAL,
                             "attention_decode_no_alloc: launch failed: %s", cudaGetErrorString(e));
      }
    }
  }
  if (previous >= 0 && previous != device_ordinal) cudaSetDevice(previous);
  return status;
}
