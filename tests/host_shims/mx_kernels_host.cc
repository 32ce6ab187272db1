// Host-side execution of mx_kernels.cu's device code (the container that wrote it has no
// GPU): CUDA keywords are shimmed, the 256 threads of a block are 256 OS threads, and
// __syncthreads() is a pthread barrier; blocks run one after another. tests/test_oracle.py
// compares the result with the oracle. Test infrastructure only.
#include <pthread.h>

#include <cmath>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

#define __global__
#define __device__
#define __forceinline__ inline
#define __restrict__
#define __shared__ static
#define __launch_bounds__(x)
struct Dim3 {
  unsigned x = 0, y = 0, z = 0;
};
static thread_local Dim3 threadIdx;
static Dim3 blockIdx;
static pthread_barrier_t block_barrier;
static inline float __uint_as_float(uint32_t u) {
  float f;
  memcpy(&f, &u, 4);
  return f;
}
// -ffp-contract=off on the command line keeps these two roundings apart, like the PTX.
static inline float __fadd_rn(float a, float b) { return a + b; }
static inline float __fmul_rn(float a, float b) { return a * b; }
static inline uint32_t __float_as_uint(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  return u;
}
#define __syncthreads() pthread_barrier_wait(&block_barrier)

#include "../../iree_b200/csrc/kernels/mx_kernels.cu"

extern "C" __attribute__((visibility("default"))) void mx_host_launch(
    const uint8_t* lhs, const uint8_t* lhs_scales, const uint8_t* rhs, const uint8_t* rhs_scales,
    float* out, int M, int N, int K, int flags) {
  pthread_barrier_init(&block_barrier, nullptr, 256);
  for (unsigned by = 0; by < (unsigned)((M + 63) / 64); ++by) {
    for (unsigned bx = 0; bx < (unsigned)((N + 63) / 64); ++bx) {
      blockIdx.x = bx;
      blockIdx.y = by;
      std::vector<std::thread> threads;
      for (unsigned t = 0; t < 256; ++t) {
        threads.emplace_back([=] {
          threadIdx.x = t;
          iree_b200_matmul_generic_mxfp4_f32(lhs, lhs_scales, rhs, rhs_scales, out, M, N, K, 1,
                                             flags);
        });
      }
      for (auto& th : threads) th.join();
    }
  }
  pthread_barrier_destroy(&block_barrier);
}
