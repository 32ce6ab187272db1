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
//   tuning      bits 0-7: 16-byte multimem requests in flight per thread (1, 2, 4 or 8);
//               bit 8: software-pipelined (the loads of
//               step i+1 are issued before the stores of step i, so the switch's reduce
//               traffic and its broadcast traffic overlap on the two link directions).
//               The block size is the launch's blockDim.x. Every rank passes the same values.
// Watchdog: off unless the table's |watchdog_cycles| is non-zero (set at attach from
// IREE_B200_COLLECTIVE_WATCHDOG_MS). A spin that exceeds it records why in the table's
// status word (host-readable: iree_b200_channel_query_collective_status) and the launch
// retires without touching more data — no trap, the context stays usable; the channel is
// then failed (its flag words are in an unknown state). Without a limit a missing peer is
// waited for indefinitely, as NCCL does.

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
  uint32_t ready;             // "range c is ready" word of fused launches
  uint32_t status;            // 0, or CollectiveConfig::kStatus* | block << 8 after a timeout
  uint64_t watchdog_cycles;   // 0 = wait forever
};
static_assert(sizeof(PeerTable) == CollectiveConfig::kTableBytes, "table layout");

enum : int { kF16 = 0, kBF16 = 1, kF32 = 2 };

// A spin outlived the watchdog: leave the reason where the host can read it; the caller
// then winds the block down.
__device__ __noinline__ void watchdog_expired(const PeerTable* table, uint32_t what) {
  atomicCAS(const_cast<uint32_t*>(&table->status), 0u, what | ((uint32_t)blockIdx.x << 8));
  __threadfence_system();
}

// Both return false when the watchdog ended the spin.
__device__ __forceinline__ bool put_flag(const PeerTable* table, uint32_t* addr) {
  uint32_t old;
  const long long start = clock64();
  const long long limit = (long long)table->watchdog_cycles;
  do {
    asm volatile("atom.global.release.sys.cas.b32 %0, [%1], 0, 1;"
                 : "=r"(old)
                 : "l"(addr)
                 : "memory");
    if (old != 0 && limit && clock64() - start > limit) {
      watchdog_expired(table, CollectiveConfig::kStatusPutStuck);
      return false;
    }
  } while (old != 0);
  return true;
}

__device__ __forceinline__ bool wait_flag(const PeerTable* table, uint32_t* addr) {
  uint32_t old;
  const long long start = clock64();
  const long long limit = (long long)table->watchdog_cycles;
  do {
    asm volatile("atom.global.acquire.sys.cas.b32 %0, [%1], 1, 0;"
                 : "=r"(old)
                 : "l"(addr)
                 : "memory");
    if (old != 1 && limit && clock64() - start > limit) {
      watchdog_expired(table, CollectiveConfig::kStatusWaitStuck);
      return false;
    }
  } while (old != 1);
  return true;
}

// Block |blockIdx.x| of every rank meets here. Callers fence (system scope) their own
// writes before, so that the release of the put covers the whole block's traffic.
// Returns false (to every thread of the block) when the watchdog gave up on a peer.
__device__ __forceinline__ bool meet_all_ranks(const PeerTable* table, int rank, int world) {
  __syncthreads();
  bool ok = true;
  if ((int)threadIdx.x < world) {
    const int peer = threadIdx.x;
    uint32_t* theirs = reinterpret_cast<uint32_t*>(table->flag_blocks[peer]) +
                       blockIdx.x * CollectiveConfig::kMaxWorld + rank;
    uint32_t* mine = reinterpret_cast<uint32_t*>(table->flag_blocks[rank]) +
                     blockIdx.x * CollectiveConfig::kMaxWorld + peer;
    ok = put_flag(table, theirs) && wait_flag(table, mine);
  }
  return __syncthreads_and(ok) != 0;
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
template <int kType, int kUnroll, bool kPipelined>
__device__ __forceinline__ void reduce_slice_nvls(uint64_t base, uint64_t begin, uint64_t end) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  const uint64_t step = stride * kUnroll;
  uint64_t i = begin + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if constexpr (!kPipelined) {
    for (; i < end; i += step) {
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
    if (i >= end) return;
    uint4 cur[kUnroll], next[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      if (i + u * stride < end) cur[u] = multimem_ld_reduce<kType>(base + (i + u * stride) * 16);
    }
    for (;;) {
      const uint64_t n = i + step;
      if (n < end) {
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
          if (n + u * stride < end) next[u] = multimem_ld_reduce<kType>(base + (n + u * stride) * 16);
        }
      }
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        if (i + u * stride < end) multimem_st(base + (i + u * stride) * 16, cur[u]);
      }
      if (n >= end) break;
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) cur[u] = next[u];
      i = n;
    }
  }
}

template <int kType, bool kNvls>
__device__ __forceinline__ void reduce_slice(const PeerTable* __restrict__ table,
                                             uint64_t multicast, uint64_t byte_offset,
                                             uint64_t vectors, int rank, int world,
                                             uint32_t tuning) {
  const uint64_t per = (vectors + world - 1) / world;
  const uint64_t begin = min(vectors, (uint64_t)rank * per);
  const uint64_t end = min(vectors, begin + per);
  if constexpr (kNvls) {
    const uint64_t base = multicast + byte_offset;
    const uint32_t unroll = tuning & 0xff;
    if (tuning & 0x100) {
      switch (unroll) {
        case 1: reduce_slice_nvls<kType, 1, true>(base, begin, end); break;
        case 2: reduce_slice_nvls<kType, 2, true>(base, begin, end); break;
        default: reduce_slice_nvls<kType, 4, true>(base, begin, end); break;
      }
    } else {
      switch (unroll) {
        case 1: reduce_slice_nvls<kType, 1, false>(base, begin, end); break;
        case 2: reduce_slice_nvls<kType, 2, false>(base, begin, end); break;
        case 8: reduce_slice_nvls<kType, 8, false>(base, begin, end); break;
        default: reduce_slice_nvls<kType, 4, false>(base, begin, end); break;
      }
    }
  } else {
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    const uint64_t first = begin + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
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

__device__ __forceinline__ bool wait_ready(const PeerTable* table, const uint32_t* ready,
                                           uint32_t value) {
  bool ok = true;
  if (threadIdx.x == 0) {
    const long long start = clock64();
    const long long limit = (long long)table->watchdog_cycles;
    for (;;) {
      uint32_t now;
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(now) : "l"(ready) : "memory");
      if ((int32_t)(now - value) >= 0) break;
      if (limit && clock64() - start > limit) {
        watchdog_expired(table, CollectiveConfig::kStatusNeverReady);
        ok = false;
        break;
      }
    }
  }
  return __syncthreads_and(ok) != 0;
}

template <int kType, bool kNvls>
__device__ __forceinline__ void all_reduce_sum_body(const PeerTable* __restrict__ table,
                                                    uint64_t multicast, uint64_t byte_offset,
                                                    uint64_t vectors, uint32_t chunks,
                                                    uint64_t chunk_stride, const uint32_t* ready,
                                                    uint32_t ready_base, int rank, int world,
                                                    uint32_t tuning) {
  for (uint32_t c = 0; c < chunks; ++c) {
    if (c > 0 && !wait_ready(table, ready, ready_base + c)) return;
    // Every rank's producer of range c has finished.
    if (!meet_all_ranks(table, rank, world)) return;
    reduce_slice<kType, kNvls>(table, multicast, byte_offset + c * chunk_stride, vectors, rank,
                               world, tuning);
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
          int rank, int world, uint32_t tuning) {                                               \
    all_reduce_sum_body<TYPE, true>(static_cast<const PeerTable*>(table), multicast,            \
                                    byte_offset, vectors, chunks, chunk_stride, ready,          \
                                    ready_base, rank, world, tuning);                           \
  }                                                                                             \
  extern "C" __global__ void __launch_bounds__(b200::CollectiveConfig::kThreads)                \
      iree_b200_all_reduce_sum_##SUFFIX##_p2p(                                                  \
          const void* table, uint64_t multicast, uint64_t byte_offset, uint64_t vectors,        \
          uint32_t chunks, uint64_t chunk_stride, const uint32_t* ready, uint32_t ready_base,   \
          int rank, int world, uint32_t tuning) {                                               \
    all_reduce_sum_body<TYPE, false>(static_cast<const PeerTable*>(table), multicast,           \
                                     byte_offset, vectors, chunks, chunk_stride, ready,         \
                                     ready_base, rank, world, tuning);                          \
  }

B200_ALL_REDUCE(f16, kF16)
B200_ALL_REDUCE(bf16, kBF16)
B200_ALL_REDUCE(f32, kF32)
