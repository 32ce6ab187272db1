// collective_kernels.cu — all-reduce(sum) over a symmetric region, executed by the GPUs
// themselves over NVLink 5 / NVSwitch instead of by NCCL.
//
// Stands behind iree_hal_command_buffer_collective (hal/command_buffer.h:300-400,
// nccl_channel.c:360-380) for the case the tensor-parallel decoder layer needs: in-place
// all_reduce(sum) of a bf16 / f16 / f32 activation that lives in a region every rank has
// mapped (include/iree_b200_hal.h, iree_b200_channel_attach_symmetric_region). Two-shot:
// rank r owns slice r of the range; it reduces that slice across all ranks and broadcasts
// the result back to all ranks.
//   *_nvls : the switch does both halves — multimem.ld_reduce pulls the slice from every
//            rank and adds in the NVSwitch (f32 accumulate), multimem.st writes it to every
//            rank's copy. Per GPU: size/W bytes in, size/W bytes out of the SMs.
//   *_p2p  : no multicast object: plain peer loads (rank order 0..W-1, f32 accumulate) and
//            peer stores.
// Ranks meet twice per launch (before: every rank's producer has finished; after: every
// slice has landed everywhere) through per-block flag words in each rank's flag block:
// put = CAS 0->1 on the peer (release.sys), wait = CAS 1->0 on our own (acquire.sys), so
// the words reset themselves and launches can follow each other without a host reset.
//
// Parameters (not the dispatch ABI — launched by the channel, not by a module):
//   table       device pointer to {bases[8], flag_blocks[8]} (this rank's mapping of every
//               rank's region / flag block)
//   multicast   multicast mapping of the region (nvls only; 0 otherwise)
//   byte_offset of the (first) range inside the region (multiple of 16)
//   vectors     number of 16-byte vectors in a range
//   chunks, chunk_stride   one launch can cover |chunks| equally sized ranges |chunk_stride|
//               bytes apart whose producers finish one after the other: range 0 is ready at
//               launch, range c once the word at |ready| has reached ready_base + c (the
//               launch layer writes it with cuStreamWriteValue32 behind producer c on the
//               dispatch stream). The ranks meet once per range and once at the very end,
//               instead of twice per launch, and there is no launch gap between ranges.
//   ready, ready_base
//   rank, world

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>

#include "collective_config.h"

namespace {

using b200::CollectiveConfig;

struct PeerTable {
  uint64_t bases[CollectiveConfig::kMaxWorld];
  uint64_t flag_blocks[CollectiveConfig::kMaxWorld];
};

enum : int { kF16 = 0, kBF16 = 1, kF32 = 2 };

constexpr long long kWatchdogCycles = 6000000000ll;  // ~3 s: trap instead of hanging the box

__device__ __forceinline__ void put_flag(uint32_t* addr) {
  uint32_t old;
  const long long start = clock64();
  do {
    asm volatile("atom.global.release.sys.cas.b32 %0, [%1], 0, 1;"
                 : "=r"(old)
                 : "l"(addr)
                 : "memory");
    if (old != 0 && clock64() - start > kWatchdogCycles) {
      printf("b200 watchdog: symmetric all-reduce put stuck (block %d thread %d)\n",
             (int)blockIdx.x, (int)threadIdx.x);
      __trap();
    }
  } while (old != 0);
}

__device__ __forceinline__ void wait_flag(uint32_t* addr) {
  uint32_t old;
  const long long start = clock64();
  do {
    asm volatile("atom.global.acquire.sys.cas.b32 %0, [%1], 1, 0;"
                 : "=r"(old)
                 : "l"(addr)
                 : "memory");
    if (old != 1 && clock64() - start > kWatchdogCycles) {
      printf("b200 watchdog: symmetric all-reduce wait stuck (block %d thread %d)\n",
             (int)blockIdx.x, (int)threadIdx.x);
      __trap();
    }
  } while (old != 1);
}

// Block |blockIdx.x| of every rank meets here. Callers fence (system scope) their own
// writes before, so that the release of the put covers the whole block's traffic.
__device__ __forceinline__ void meet_all_ranks(const PeerTable* table, int rank, int world) {
  __syncthreads();
  if ((int)threadIdx.x < world) {
    const int peer = threadIdx.x;
    uint32_t* theirs = reinterpret_cast<uint32_t*>(table->flag_blocks[peer]) +
                       blockIdx.x * CollectiveConfig::kMaxWorld + rank;
    uint32_t* mine = reinterpret_cast<uint32_t*>(table->flag_blocks[rank]) +
                     blockIdx.x * CollectiveConfig::kMaxWorld + peer;
    put_flag(theirs);
    wait_flag(mine);
  }
  __syncthreads();
}

template <int kType>
__device__ __forceinline__ uint4 multimem_ld_reduce(uint64_t addr) {
  uint4 v;
  if constexpr (kType == kBF16) {
    asm volatile(
        "multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
        : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
        : "l"(addr)
        : "memory");
  } else if constexpr (kType == kF16) {
    asm volatile(
        "multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0, %1, %2, %3}, [%4];"
        : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
        : "l"(addr)
        : "memory");
  } else {
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(addr)
                 : "memory");
  }
  return v;
}

__device__ __forceinline__ void multimem_st(uint64_t addr, const uint4& v) {
  // The payload is moved as four 32-bit words whatever the element type.
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr),
               "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

__device__ __forceinline__ uint4 ld_peer(uint64_t addr) {
  uint4 v;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(addr)
               : "memory");
  return v;
}

__device__ __forceinline__ void st_peer(uint64_t addr, const uint4& v) {
  asm volatile("st.relaxed.sys.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "r"(v.x),
               "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

template <int kType>
__device__ __forceinline__ void accumulate(float (&acc)[8], const uint4& v) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
  if constexpr (kType == kF32) {
#pragma unroll
    for (int e = 0; e < 4; ++e) acc[e] += __uint_as_float(w[e]);
  } else if constexpr (kType == kBF16) {
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      acc[2 * e] += __uint_as_float(w[e] << 16);
      acc[2 * e + 1] += __uint_as_float(w[e] & 0xffff0000u);
    }
  } else {
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const __half2 h = *reinterpret_cast<const __half2*>(&w[e]);
      acc[2 * e] += __low2float(h);
      acc[2 * e + 1] += __high2float(h);
    }
  }
}

template <int kType>
__device__ __forceinline__ uint4 round_result(const float (&acc)[8]) {
  uint32_t w[4];
  if constexpr (kType == kF32) {
#pragma unroll
    for (int e = 0; e < 4; ++e) w[e] = __float_as_uint(acc[e]);
  } else if constexpr (kType == kBF16) {
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const __nv_bfloat162 p = __floats2bfloat162_rn(acc[2 * e], acc[2 * e + 1]);
      w[e] = *reinterpret_cast<const uint32_t*>(&p);
    }
  } else {
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const __half2 p = __floats2half2_rn(acc[2 * e], acc[2 * e + 1]);
      w[e] = *reinterpret_cast<const uint32_t*>(&p);
    }
  }
  return make_uint4(w[0], w[1], w[2], w[3]);
}

// Slice of rank r: vectors [r*per, min(vectors, (r+1)*per)), per = ceil(vectors / world).
template <int kType, bool kNvls>
__device__ __forceinline__ void reduce_slice(const PeerTable* __restrict__ table,
                                             uint64_t multicast, uint64_t byte_offset,
                                             uint64_t vectors, int rank, int world) {
  constexpr int kUnroll = CollectiveConfig::kUnroll;
  const uint64_t per = (vectors + world - 1) / world;
  const uint64_t begin = min(vectors, (uint64_t)rank * per);
  const uint64_t end = min(vectors, begin + per);
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  const uint64_t first = begin + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if constexpr (kNvls) {
    const uint64_t base = multicast + byte_offset;
    for (uint64_t i = first; i < end; i += stride * kUnroll) {
      uint4 v[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        if (i + u * stride < end) v[u] = multimem_ld_reduce<kType>(base + (i + u * stride) * 16);
      }
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        if (i + u * stride < end) multimem_st(base + (i + u * stride) * 16, v[u]);
      }
    }
  } else {
    for (uint64_t i = first; i < end; i += stride) {
      float acc[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] = 0.f;
      uint4 v[CollectiveConfig::kMaxWorld];
#pragma unroll
      for (int p = 0; p < CollectiveConfig::kMaxWorld; ++p) {
        if (p < world) v[p] = ld_peer(table->bases[p] + byte_offset + i * 16);
      }
#pragma unroll
      for (int p = 0; p < CollectiveConfig::kMaxWorld; ++p) {
        if (p < world) accumulate<kType>(acc, v[p]);
      }
      const uint4 r = round_result<kType>(acc);
#pragma unroll
      for (int p = 0; p < CollectiveConfig::kMaxWorld; ++p) {
        if (p < world) st_peer(table->bases[p] + byte_offset + i * 16, r);
      }
    }
  }
}

__device__ __forceinline__ void wait_ready(const uint32_t* ready, uint32_t value) {
  if (threadIdx.x == 0) {
    const long long start = clock64();
    for (;;) {
      uint32_t now;
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(now) : "l"(ready) : "memory");
      if ((int32_t)(now - value) >= 0) break;
      if (clock64() - start > kWatchdogCycles) {
        printf("b200 watchdog: all-reduce range never became ready (block %d, want %u, have %u)\n",
               (int)blockIdx.x, value, now);
        __trap();
      }
    }
  }
  __syncthreads();
}

template <int kType, bool kNvls>
__device__ __forceinline__ void all_reduce_sum_body(const PeerTable* __restrict__ table,
                                                    uint64_t multicast, uint64_t byte_offset,
                                                    uint64_t vectors, uint32_t chunks,
                                                    uint64_t chunk_stride, const uint32_t* ready,
                                                    uint32_t ready_base, int rank, int world) {
  for (uint32_t c = 0; c < chunks; ++c) {
    if (c > 0) wait_ready(ready, ready_base + c);
    // Every rank's producer of range c has finished.
    meet_all_ranks(table, rank, world);
    reduce_slice<kType, kNvls>(table, multicast, byte_offset + c * chunk_stride, vectors, rank,
                               world);
  }
  // Every slice of every range has landed everywhere.
  __threadfence_system();
  meet_all_ranks(table, rank, world);
}

}  // namespace

#define B200_ALL_REDUCE(SUFFIX, TYPE)                                                           \
  extern "C" __global__ void __launch_bounds__(b200::CollectiveConfig::kThreads)                \
      iree_b200_all_reduce_sum_##SUFFIX##_nvls(                                                 \
          const void* table, uint64_t multicast, uint64_t byte_offset, uint64_t vectors,        \
          uint32_t chunks, uint64_t chunk_stride, const uint32_t* ready, uint32_t ready_base,   \
          int rank, int world) {                                                                \
    all_reduce_sum_body<TYPE, true>(static_cast<const PeerTable*>(table), multicast,            \
                                    byte_offset, vectors, chunks, chunk_stride, ready,          \
                                    ready_base, rank, world);                                   \
  }                                                                                             \
  extern "C" __global__ void __launch_bounds__(b200::CollectiveConfig::kThreads)                \
      iree_b200_all_reduce_sum_##SUFFIX##_p2p(                                                  \
          const void* table, uint64_t multicast, uint64_t byte_offset, uint64_t vectors,        \
          uint32_t chunks, uint64_t chunk_stride, const uint32_t* ready, uint32_t ready_base,   \
          int rank, int world) {                                                                \
    all_reduce_sum_body<TYPE, false>(static_cast<const PeerTable*>(table), multicast,           \
                                     byte_offset, vectors, chunks, chunk_stride, ready,         \
                                     ready_base, rank, world);                                  \
  }

B200_ALL_REDUCE(f16, kF16)
B200_ALL_REDUCE(bf16, kBF16)
B200_ALL_REDUCE(f32, kF32)
