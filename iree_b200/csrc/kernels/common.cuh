// common.cuh — sm_100a PTX wrappers shared by the dispatch kernels.
//
// Hand-written inline PTX for: mbarrier, TMA (cp.async.bulk.tensor), tcgen05
// (alloc / mma / commit / ld / st / fences), clusters. No CUTLASS/CuTe types.
//
// All kernels in this directory follow IREE's custom-dispatch CUDA ABI
// (reference: samples/custom_dispatch/cuda/kernels/kernels.cu:7-34): extern "C",
// one pointer per binding in layout order, then one 32-bit value per push constant.
// Kernels that are fed by TMA additionally take `const __grid_constant__ TensorMap`
// parameters AFTER the ABI parameters; the thin HAL layer encodes and appends them at
// dispatch time from the `<kernel>__tma` recipe global (see tma_recipe.h).

#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>

#include "gemm_config.h"

namespace b200 {

// 128-byte opaque TMA descriptor (same size/alignment as CUtensorMap).
struct alignas(64) TensorMap {
  uint64_t opaque[16];
};

#define B200_DEVICE __device__ __forceinline__

// Spin-wait watchdog: a mis-programmed barrier must surface as a CUDA error, never
// as a hung GPU. ~4 s at 1.9 GHz before trapping.
#ifndef B200_WATCHDOG_CYCLES
#define B200_WATCHDOG_CYCLES (8000000000ll)
#endif

B200_DEVICE uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

B200_DEVICE uint32_t lane_id() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%laneid;" : "=r"(r));
  return r;
}

B200_DEVICE uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}

B200_DEVICE uint32_t cluster_id_x() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%clusterid.x;" : "=r"(r));
  return r;
}

B200_DEVICE uint32_t num_clusters_x() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%nclusterid.x;" : "=r"(r));
  return r;
}

// Returns true in exactly one lane of a fully converged warp.
B200_DEVICE bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t"
      "}"
      : "=r"(pred));
  return pred != 0;
}

B200_DEVICE void cluster_arrive() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
}
B200_DEVICE void cluster_wait() {
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
B200_DEVICE void cluster_sync() {
  cluster_arrive();
  cluster_wait();
}

// shared::cta address -> shared::cluster address of the same offset in CTA `rank`.
B200_DEVICE uint32_t map_to_cta(uint32_t smem_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(rank));
  return r;
}

//===----------------------------------------------------------------------===//
// mbarrier
//===----------------------------------------------------------------------===//

B200_DEVICE void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}

// Makes mbarrier inits visible to the async proxy and to peer CTAs.
B200_DEVICE void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

B200_DEVICE void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

B200_DEVICE void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
               : "memory");
}

B200_DEVICE void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}

// Arrive on a barrier given by a shared::cluster address (possibly in a peer CTA).
// Default semantics (.release.cta): an explicit .release.cluster makes ptxas emit
// MEMBAR.ALL.GPU + ERRBAR in front of every arrive, which costs microseconds inside a
// pipeline loop. The data these arrives order (TMEM reads, TMA-written smem) is ordered
// by tcgen05 fences / the async proxy, not by generic-proxy release semantics.
B200_DEVICE void mbar_arrive_cluster(uint32_t cluster_bar) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_bar) : "memory");
}

B200_DEVICE bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t"
      "}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}

B200_DEVICE void mbar_wait(uint32_t bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  long long start = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (clock64() - start > B200_WATCHDOG_CYCLES) {
      printf("b200 watchdog: mbarrier 0x%x parity %u stuck (block %d thread %d)\n", bar, parity,
             (int)blockIdx.x, (int)threadIdx.x);
      __trap();
    }
  }
}

//===----------------------------------------------------------------------===//
// TMA
//===----------------------------------------------------------------------===//

B200_DEVICE void tma_prefetch_desc(const TensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(map)) : "memory");
}

// ---- Tensor maps built on the device (kernels launched without trailing CUtensorMap
// parameters, i.e. by an unmodified IREE CUDA driver) -----------------------------------------
// A 128-byte template (encoded once on the host for a dummy shape) is copied to shared memory,
// patched with tensormap.replace, and published to a global slot with the fused copy +
// release fence; whoever then issues TMA through the slot first runs the acquire fence.

// One warp, converged. |staging| (shared, 128-byte aligned) and |slot| (global, 128-byte
// aligned) hold one map; rank 3, dense tensor: strides follow from the extents.
B200_DEVICE void tensormap_build_rank3(TensorMap* slot, TensorMap* staging, const uint64_t* tmpl,
                                       const void* base, uint32_t dim0, uint32_t dim1, uint32_t dim2,
                                       uint32_t elem_size) {
  const uint32_t lane = threadIdx.x & 31;
  reinterpret_cast<uint32_t*>(staging)[lane] = reinterpret_cast<const uint32_t*>(tmpl)[lane];
  __syncwarp();
  if (lane == 0) {
    const uint64_t s = static_cast<uint64_t>(smem_u32(staging));
    const uint64_t stride0 = (uint64_t)dim0 * elem_size, stride1 = stride0 * dim1;
    asm volatile("tensormap.replace.tile.global_address.shared::cta.b1024.b64 [%0], %1;" ::"l"(s),
                 "l"(reinterpret_cast<uint64_t>(base))
                 : "memory");
    asm volatile("tensormap.replace.tile.global_dim.shared::cta.b1024.b32 [%0], 0, %1;" ::"l"(s),
                 "r"(dim0)
                 : "memory");
    asm volatile("tensormap.replace.tile.global_dim.shared::cta.b1024.b32 [%0], 1, %1;" ::"l"(s),
                 "r"(dim1)
                 : "memory");
    asm volatile("tensormap.replace.tile.global_dim.shared::cta.b1024.b32 [%0], 2, %1;" ::"l"(s),
                 "r"(dim2)
                 : "memory");
    // Byte strides of dimensions 1 and 2 (multiples of 16; dimension 0 is dense).
    asm volatile("tensormap.replace.tile.global_stride.shared::cta.b1024.b64 [%0], 0, %1;" ::"l"(s),
                 "l"(stride0)
                 : "memory");
    asm volatile("tensormap.replace.tile.global_stride.shared::cta.b1024.b64 [%0], 1, %1;" ::"l"(s),
                 "l"(stride1)
                 : "memory");
  }
  __syncwarp();
  asm volatile(
      "tensormap.cp_fenceproxy.global.shared::cta.tensormap::generic.release.gpu.sync.aligned"
      " [%0], [%1], 128;" ::"l"(reinterpret_cast<uint64_t>(slot)),
      "r"(smem_u32(staging))
      : "memory");
}

// Slot index of the calling CTA: its SM. The kernels that build maps need more than half an
// SM's shared memory, so at most one of their CTAs is resident per SM and a slot has one
// writer / reader at a time — also across back-to-back or overlapping launches.
constexpr int kTensorMapSlotSms = 512;
B200_DEVICE uint32_t tensormap_slot_index() {
  uint32_t smid;
  asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
  return smid % kTensorMapSlotSms;
}

B200_DEVICE void tensormap_fence_acquire(const TensorMap* slot) {
  asm volatile("fence.proxy.tensormap::generic.acquire.gpu [%0], 128;" ::"l"(
                   reinterpret_cast<uint64_t>(slot))
               : "memory");
}

// 3-D tiled load into this CTA's shared memory; completes `bytes` on `bar`.
// kCtaGroup == 2: `bar` is a shared::cluster address (the pair leader's barrier).
template <int kCtaGroup>
B200_DEVICE void tma_load_3d(uint32_t dst_smem, const TensorMap* map, uint32_t bar, int c0, int c1,
                             int c2) {
  if constexpr (kCtaGroup == 1) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(dst_smem),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
  } else {
    asm volatile(
        "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(dst_smem),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
  }
}

// 4-D form of tma_load_3d (tiled operands: linalg.mmt4d).
template <int kCtaGroup>
B200_DEVICE void tma_load_4d(uint32_t dst_smem, const TensorMap* map, uint32_t bar, int c0, int c1,
                             int c2, int c3) {
  if constexpr (kCtaGroup == 1) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(dst_smem),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
  } else {
    asm volatile(
        "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(dst_smem),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
  }
}

B200_DEVICE void tma_store_4d(const TensorMap* map, uint32_t src_smem, int c0, int c1, int c2,
                              int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(
          reinterpret_cast<uint64_t>(map)),
      "r"(src_smem), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}

// 1-D bulk copy global -> this CTA's shared memory (no tensor map: plain pointers, so the
// kernel stays launchable through the unmodified custom-dispatch ABI). |bytes| and both
// addresses are multiples of 16; completes |bytes| on |bar|.
B200_DEVICE void bulk_load_1d(uint32_t dst_smem, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
          "r"(dst_smem),
      "l"(src), "r"(bytes), "r"(bar)
      : "memory");
}

B200_DEVICE void st_shared_cluster_f32(uint32_t cluster_addr, float v) {
  asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(cluster_addr), "f"(v) : "memory");
}

B200_DEVICE float ld_shared_cluster_f32(uint32_t cluster_addr) {
  float v;
  asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(cluster_addr) : "memory");
  return v;
}

B200_DEVICE uint64_t l2_policy_evict_last() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
}
B200_DEVICE uint64_t l2_policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}

// tma_load_3d with an L2 eviction-priority hint.
template <int kCtaGroup>
B200_DEVICE void tma_load_3d_hint(uint32_t dst_smem, const TensorMap* map, uint32_t bar, int c0,
                                  int c1, int c2, uint64_t policy) {
  if constexpr (kCtaGroup == 1) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
        " [%0], [%1, {%3, %4, %5}], [%2], %6;" ::"r"(dst_smem),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "l"(policy)
        : "memory");
  } else {
    asm volatile(
        "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
        ".L2::cache_hint [%0], [%1, {%3, %4, %5}], [%2], %6;" ::"r"(dst_smem),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "l"(policy)
        : "memory");
  }
}

B200_DEVICE void tma_store_3d(const TensorMap* map, uint32_t src_smem, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(
          reinterpret_cast<uint64_t>(map)),
      "r"(src_smem), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}

B200_DEVICE void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }

template <int kPending>
B200_DEVICE void tma_store_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(kPending) : "memory");
}
template <int kPending>
B200_DEVICE void tma_store_wait() {
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(kPending) : "memory");
}

//===----------------------------------------------------------------------===//
// tcgen05: TMEM allocation, MMA, commit, ld/st, fences
//===----------------------------------------------------------------------===//

template <int kCtaGroup>
B200_DEVICE void tmem_alloc(uint32_t dst_smem, uint32_t columns) {
  if constexpr (kCtaGroup == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem),
                 "r"(columns)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  } else {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem),
                 "r"(columns)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
}

template <int kCtaGroup>
B200_DEVICE void tmem_dealloc(uint32_t tmem_addr, uint32_t columns) {
  if constexpr (kCtaGroup == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_addr),
                 "r"(columns)
                 : "memory");
  } else {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_addr),
                 "r"(columns)
                 : "memory");
  }
}

B200_DEVICE void tcgen05_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
B200_DEVICE void tcgen05_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}


// D[tmem] (+)= A[smem desc] * B[smem desc]. Issued by ONE thread.
template <int kCtaGroup, int kKind>
B200_DEVICE void umma_ss(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                         uint32_t accumulate) {
  if constexpr (kCtaGroup == 1 && kKind == KIND_F16) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
  } else if constexpr (kCtaGroup == 2 && kKind == KIND_F16) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
  } else if constexpr (kCtaGroup == 1 && kKind == KIND_I8) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
  } else if constexpr (kCtaGroup == 2 && kKind == KIND_I8) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
  } else if constexpr (kCtaGroup == 1 && kKind == KIND_F8) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
  } else {
    static_assert(kKind == KIND_F8 && kCtaGroup == 2, "unhandled MMA kind");
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
  }
}

// D[tmem] (+)= A[tmem] * B[smem desc]  (A operand read from tensor memory).
template <int kCtaGroup>
B200_DEVICE void umma_ts_f16(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc,
                             uint32_t accumulate) {
  if constexpr (kCtaGroup == 1) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem_d),
        "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
  } else {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem_d),
        "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
  }
}

// Arrives on `bar` (shared::cta address in this CTA) once all previously issued
// tcgen05.mma of this thread have completed. Implies fence::before_thread_sync.
B200_DEVICE void umma_commit_1cta(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                   bar)
               : "memory");
}

// cta_group::2 form: arrives on the barrier at the same offset in every CTA of `mask`.
B200_DEVICE void umma_commit_2cta(uint32_t bar, uint16_t mask) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64"
      " [%0], %1;" ::"r"(bar),
      "h"(mask)
      : "memory");
}

template <int kCtaGroup>
B200_DEVICE void umma_commit(uint32_t bar) {
  if constexpr (kCtaGroup == 1) {
    umma_commit_1cta(bar);
  } else {
    umma_commit_2cta(bar, 0x3);
  }
}

// TMEM -> registers: this warp's 32 lanes x 32 consecutive 32-bit columns.
B200_DEVICE void tmem_ld_32x32b_x32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
        "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
        "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]),
        "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
        "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}

B200_DEVICE void tmem_ld_32x32b_x16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
        "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
        "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}

B200_DEVICE void tmem_st_32x32b_x16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
      "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}

B200_DEVICE void tmem_st_32x32b_x32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(
          taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
      "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]),
      "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]),
      "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]),
      "r"(r[30]), "r"(r[31])
      : "memory");
}

B200_DEVICE void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
B200_DEVICE void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

//===----------------------------------------------------------------------===//
// UMMA descriptors (bit layouts: PTX ISA "tcgen05 matrix descriptor" /
// "instruction descriptor"; cross-checked against the public CuTe definitions).
//===----------------------------------------------------------------------===//

// Shared-memory matrix descriptor, SWIZZLE_128B:
//   [0,14)  start address >> 4      [16,30) leading-dim byte offset >> 4
//   [32,46) stride-dim byte offset >> 4   [46,48) version = 1   [61,64) layout = 2
B200_DEVICE uint64_t make_smem_desc_sw128(uint32_t smem_addr, uint32_t lbo_bytes,
                                          uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= 1ull << 46;
  d |= 2ull << 61;
  return d;
}

// Unswizzled K-major operand: core matrices of 8 rows x 16 bytes stored contiguously
// (128 bytes); |lbo_bytes| between core matrices adjacent in K, |sbo_bytes| between 8-row
// groups (layout type 0).
B200_DEVICE uint64_t make_smem_desc_interleaved(uint32_t smem_addr, uint32_t lbo_bytes,
                                                uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= 1ull << 46;
  return d;
}

B200_DEVICE void named_barrier_sync(uint32_t id, uint32_t threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

}  // namespace b200
