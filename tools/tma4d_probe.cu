// Stand-alone probe of rank-4 tensor-map loads / stores (the tile-direct mmt4d maps of
// gemm_sm100.cuh): one CTA, one box, the shared-memory image copied out and compared on the
// host. Each case runs in its own process (a fault poisons the context).
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -o iree_b200/_build/tma4d_probe \
//        tools/tma4d_probe.cu -lcuda
//   tma4d_probe <load|store> <dense|permuted> <swizzle 0|3> <k0 bytes> <k1 box> <r0> <r1 box>
//
// Tensor: [R1 = 8][K1 = 16][R0][K0] of 16-bit elements (K0 * 2 = k0 bytes).
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#define CHECK(x)                                                         \
  do {                                                                   \
    auto e_ = (x);                                                       \
    if ((int)e_ != 0) {                                                  \
      printf("FAILED %s -> %d (line %d)\n", #x, (int)e_, __LINE__);      \
      return 1;                                                          \
    }                                                                    \
  } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}

__global__ void probe(const __grid_constant__ CUtensorMap map, uint16_t* out, const uint16_t* image,
                      int bytes, int store, int c1, int c3) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(8) uint64_t bar;
  const uint32_t tile = (smem_u32(smem) + 1023u) & ~1023u;
  uint8_t* tile_p = smem + (tile - smem_u32(smem));
  if (store) {
    for (int i = threadIdx.x; i < bytes / 2; i += blockDim.x) {
      reinterpret_cast<uint16_t*>(tile_p)[i] = image[i];
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (threadIdx.x == 0) {
      asm volatile(
          "cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(
              reinterpret_cast<uint64_t>(&map)),
          "r"(tile), "r"(0), "r"(c1), "r"(0), "r"(c3)
          : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
    return;
  }
  const uint32_t b = smem_u32(&bar);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(b));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(bytes)
                 : "memory");
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(tile),
        "l"(reinterpret_cast<uint64_t>(&map)), "r"(b), "r"(0), "r"(c1), "r"(0), "r"(c3)
        : "memory");
  }
  uint32_t done = 0;
  while (!done) {
    asm volatile(
        "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\nselp.u32 %0, 1, 0, p;\n}"
        : "=r"(done)
        : "r"(b)
        : "memory");
  }
  for (int i = threadIdx.x; i < bytes / 2; i += blockDim.x) {
    out[i] = reinterpret_cast<uint16_t*>(tile_p)[i];
  }
}

int main(int argc, char** argv) {
  if (argc < 8) return 2;
  const bool store = !strcmp(argv[1], "store");
  const bool permuted = !strcmp(argv[2], "permuted");
  const int swizzle = atoi(argv[3]);
  const int K0 = atoi(argv[4]) / 2, bk1 = atoi(argv[5]), R0 = atoi(argv[6]), br1 = atoi(argv[7]);
  const int R1 = 8, K1 = 16;
  const size_t elems = (size_t)R1 * K1 * R0 * K0;
  CHECK(cuInit(0));
  CHECK(cudaSetDevice(0));
  std::vector<uint16_t> host(elems);
  for (size_t i = 0; i < elems; ++i) host[i] = (uint16_t)i;
  uint16_t *tensor = nullptr, *out = nullptr, *image = nullptr;
  const int box_elems = K0 * bk1 * R0 * br1, bytes = box_elems * 2;
  CHECK(cudaMalloc(&tensor, elems * 2));
  CHECK(cudaMalloc(&out, bytes));
  CHECK(cudaMalloc(&image, bytes));
  CHECK(cudaMemset(out, 0xff, bytes));
  if (store) {
    CHECK(cudaMemset(tensor, 0xff, elems * 2));
  } else {
    CHECK(cudaMemcpy(tensor, host.data(), elems * 2, cudaMemcpyHostToDevice));
  }
  std::vector<uint16_t> img(box_elems);
  for (int i = 0; i < box_elems; ++i) img[i] = (uint16_t)(i + 1);
  CHECK(cudaMemcpy(image, img.data(), bytes, cudaMemcpyHostToDevice));

  // Map dims: permuted = (k0, k1, r0, r1), dense = (k0, r0, k1, r1).
  cuuint64_t dims[4], strides[3];
  cuuint32_t box[4], es[4] = {1, 1, 1, 1};
  const cuuint64_t s_k0 = 2, s_r0 = (cuuint64_t)K0 * 2, s_k1 = s_r0 * R0, s_r1 = s_k1 * K1;
  (void)s_k0;
  if (permuted) {
    dims[0] = K0, dims[1] = K1, dims[2] = R0, dims[3] = R1;
    strides[0] = s_k1, strides[1] = s_r0, strides[2] = s_r1;
    box[0] = K0, box[1] = bk1, box[2] = R0, box[3] = br1;
  } else {
    dims[0] = K0, dims[1] = R0, dims[2] = K1, dims[3] = R1;
    strides[0] = s_r0, strides[1] = s_k1, strides[2] = s_r1;
    box[0] = K0, box[1] = R0, box[2] = bk1, box[3] = br1;
  }
  CUtensorMap map;
  CHECK(cuTensorMapEncodeTiled(&map, CU_TENSOR_MAP_DATA_TYPE_UINT16, 4, tensor, dims, strides, box, es,
                               CU_TENSOR_MAP_INTERLEAVE_NONE, (CUtensorMapSwizzle)swizzle,
                               CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE));
  const int smem_bytes = bytes + 2048;
  CHECK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
  probe<<<1, 128, smem_bytes>>>(map, out, image, bytes, store ? 1 : 0, /*c1=*/permuted ? 8 : 0,
                                /*c3=*/0);
  // dense order: coordinate 2 is k1; keep the box at the origin there (c1 is r0's slot).
  CHECK(cudaDeviceSynchronize());
  if (store) {
    std::vector<uint16_t> back(elems);
    CHECK(cudaMemcpy(back.data(), tensor, elems * 2, cudaMemcpyDeviceToHost));
    size_t written = 0;
    for (size_t i = 0; i < elems; ++i) written += back[i] != 0xffff;
    printf("store ok: %zu of %zu elements written (box %d)\n", written, elems, box_elems);
    return 0;
  }
  std::vector<uint16_t> got(box_elems);
  CHECK(cudaMemcpy(got.data(), out, bytes, cudaMemcpyDeviceToHost));
  // Expected unswizzled image: box element (b3, b2, b1, b0) at ((b3 * box2 + b2) * box1 + b1) * box0 + b0.
  size_t bad = 0;
  for (int b3 = 0; b3 < (int)box[3]; ++b3)
    for (int b2 = 0; b2 < (int)box[2]; ++b2)
      for (int b1 = 0; b1 < (int)box[1]; ++b1)
        for (int b0 = 0; b0 < (int)box[0]; ++b0) {
          size_t at = (((size_t)b3 * box[2] + b2) * box[1] + b1) * box[0] + b0;
          if (swizzle == 3) {
            size_t byte = at * 2;
            byte ^= ((byte >> 7) & 7) << 4;
            at = byte / 2;
          }
          const int c1 = permuted ? 8 : 0;
          size_t src;
          if (permuted) {
            src = (size_t)b3 * (s_r1 / 2) + (size_t)(b1 + c1) * (s_k1 / 2) + (size_t)b2 * (s_r0 / 2) + b0;
          } else {
            src = (size_t)b3 * (s_r1 / 2) + (size_t)b2 * (s_k1 / 2) + (size_t)b1 * (s_r0 / 2) + b0;
          }
          bad += got[at] != host[src];
        }
  printf("load ok: %zu mismatches of %d\n", bad, box_elems);
  return 0;
}
