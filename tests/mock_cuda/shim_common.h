// shim_common.h — compiles the SIMT kernels of iree_b200/csrc/kernels for the HOST: CUDA keywords
// become no-ops, the 256 threads of a block are OS threads, __syncthreads() is a pthread
// barrier. Used by mock_kernels_*.cc (the mock driver runs these when a kernel is launched)
// and by tests/host_shims. Test infrastructure only.
#pragma once
#include <pthread.h>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#undef __global__
#undef __device__
#undef __host__
#undef __forceinline__
#undef __restrict__
#undef __shared__
#undef __launch_bounds__
#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __restrict__
#define __shared__ static
#define __launch_bounds__(...)
namespace {
struct Dim3 { unsigned x = 0, y = 0, z = 0; };
thread_local Dim3 threadIdx;
Dim3 blockIdx, gridDim;
pthread_barrier_t block_barrier;
inline float __uint_as_float(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
inline uint32_t __float_as_uint(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
inline float __fadd_rn(float a, float b) { return a + b; }
inline float __fmul_rn(float a, float b) { return a * b; }
}
#define __syncthreads() pthread_barrier_wait(&block_barrier)

namespace {
// Runs |body| for every thread of every block of the grid; blocks run one after another.
template <class F>
void run_grid(const unsigned grid[3], F body) {
  pthread_barrier_init(&block_barrier, nullptr, 256);
  gridDim.x = grid[0];
  gridDim.y = grid[1];
  gridDim.z = grid[2];
  for (unsigned z = 0; z < grid[2]; ++z) {
    for (unsigned y = 0; y < grid[1]; ++y) {
      for (unsigned x = 0; x < grid[0]; ++x) {
        blockIdx.x = x;
        blockIdx.y = y;
        blockIdx.z = z;
        std::vector<std::thread> threads;
        for (unsigned t = 0; t < 256; ++t) {
          threads.emplace_back([&body, t] {
            threadIdx.x = t;
            body();
          });
        }
        for (auto& th : threads) th.join();
      }
    }
  }
  pthread_barrier_destroy(&block_barrier);
}
template <class T>
T arg(void** params, int i) {
  T v;
  memcpy(&v, params[i], sizeof(T));
  return v;
}
}  // namespace
