// mock_cuda.cc — a stand-in libcuda.so.1 for exercising the launch layer's HOST logic on a
// machine without a GPU (tests/test_hal_mock_cpu.py builds it and puts it on LD_LIBRARY_PATH
// of a pytest subprocess; libiree_b200.so dlopen()s "libcuda.so.1" and finds this).
//
// TEST INFRASTRUCTURE ONLY — never shipped, never on the product path. It does not compute:
//   * memory is host memory ("device pointers" are host addresses), copies / memsets run
//     immediately, streams and events are ordering-free handles (everything is synchronous);
//   * modules keep a copy of their image; cuModuleGetGlobal finds `__device__ const` globals
//     (the `<kernel>__tma` recipes) in the cubin's ELF symbol table, so recipe evaluation and
//     cuTensorMapEncodeTiled's arguments are real;
//   * cuLaunchKernel appends (kernel name, grid, block, shared memory) to a log the tests read
//     back through mock_cuda_take_launches(). If the library named by MOCK_CUDA_KERNELS
//     exports mock_run_<kernel> (mock_kernels_*.cc: the SIMT kernels' own source compiled for
//     the host, one OS thread per CUDA thread) the launch also RUNS, with the parameter array
//     exactly as the GPU would get it; the tcgen05 / TMA kernels have no host form and are
//     only logged — their results are what the GPU suite is for;
//   * graphs record their nodes and replay them in insertion order (dependencies always point
//     backwards in the launch layer), so fills / copies / updates in graph mode are checked.
// Reports one device: compute capability 10.0, 148 SMs, 227 KiB opt-in shared memory.

#include <cuda.h>
#include <dlfcn.h>
#include <elf.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <mutex>
#include <string>
#include <vector>

#define MOCK_API extern "C" __attribute__((visibility("default")))

namespace {

struct Module {
  std::vector<uint8_t> image;
};
struct Function {
  std::string name;
};
struct Graph {
  std::vector<std::function<void()>> nodes;
};
struct Launch {
  std::string name;
  unsigned grid[3], block[3], shared;
  uint64_t args[4];  // first four parameter slots (matmul dispatches only, else 0)
};

std::mutex g_mutex;
std::vector<Launch> g_launches;
struct EncodedMap {
  uint32_t type, rank, swizzle;
  uint64_t base, dims[5], strides[5];
  uint32_t box[5];
};
std::vector<EncodedMap> g_maps;
int g_dummy_ctx, g_dummy_stream_storage;

using RunFn = int (*)(const unsigned grid[3], const unsigned block[3], void** params);

void* kernels_library() {
  static void* lib = [] {
    const char* path = getenv("MOCK_CUDA_KERNELS");
    return path ? dlopen(path, RTLD_NOW | RTLD_LOCAL) : nullptr;
  }();
  return lib;
}

CUresult record_launch(const Function* f, unsigned gx, unsigned gy, unsigned gz, unsigned bx,
                       unsigned by, unsigned bz, unsigned shared, void** params) {
  {
    Launch l = {f->name, {gx, gy, gz}, {bx, by, bz}, shared, {0, 0, 0, 0}};
    // Dispatch parameters are 8-byte slots (command_buffer.cc record_dispatch); every matmul
    // export has at least four (three bindings + M).
    if (f->name.rfind("iree_b200_matmul", 0) == 0) {
      for (int i = 0; i < 4; ++i) memcpy(&l.args[i], params[i], sizeof(uint64_t));
    }
    std::lock_guard<std::mutex> lock(g_mutex);
    g_launches.push_back(l);
  }
  if (void* lib = kernels_library()) {
    if (auto run = reinterpret_cast<RunFn>(dlsym(lib, ("mock_run_" + f->name).c_str()))) {
      const unsigned grid[3] = {gx, gy, gz}, block[3] = {bx, by, bz};
      if (run(grid, block, params) != 0) return CUDA_ERROR_LAUNCH_FAILED;
    }
  }
  return CUDA_SUCCESS;
}

void do_memset(CUdeviceptr dst, unsigned value, size_t element_size, size_t count) {
  uint8_t* p = reinterpret_cast<uint8_t*>(dst);
  for (size_t i = 0; i < count; ++i) memcpy(p + i * element_size, &value, element_size);
}

// Address and size of ELF symbol |name| inside a cubin image (ELF64, little endian).
bool find_symbol(std::vector<uint8_t>& image, const char* name, uint8_t** out, size_t* out_size) {
  if (image.size() < sizeof(Elf64_Ehdr) || memcmp(image.data(), ELFMAG, SELFMAG) != 0) return false;
  const auto* eh = reinterpret_cast<const Elf64_Ehdr*>(image.data());
  if (eh->e_shoff == 0 || eh->e_shoff + (uint64_t)eh->e_shnum * sizeof(Elf64_Shdr) > image.size()) {
    return false;
  }
  const auto* sh = reinterpret_cast<const Elf64_Shdr*>(image.data() + eh->e_shoff);
  for (int i = 0; i < eh->e_shnum; ++i) {
    if (sh[i].sh_type != SHT_SYMTAB) continue;
    const Elf64_Shdr& strtab = sh[sh[i].sh_link];
    const auto* syms = reinterpret_cast<const Elf64_Sym*>(image.data() + sh[i].sh_offset);
    const size_t n = sh[i].sh_size / sizeof(Elf64_Sym);
    for (size_t s = 0; s < n; ++s) {
      const char* sym_name =
          reinterpret_cast<const char*>(image.data() + strtab.sh_offset + syms[s].st_name);
      if (strcmp(sym_name, name) != 0 || syms[s].st_shndx >= eh->e_shnum) continue;
      const Elf64_Shdr& sec = sh[syms[s].st_shndx];
      if (sec.sh_type == SHT_NOBITS) return false;
      if (sec.sh_offset + syms[s].st_value + syms[s].st_size > image.size()) return false;
      *out = image.data() + sec.sh_offset + syms[s].st_value;
      *out_size = syms[s].st_size;
      return true;
    }
  }
  return false;
}

}  // namespace

// ---- test-side accessors -------------------------------------------------------------------

// Writes up to |capacity| bytes of "name gx gy gz bx by bz shared a0 a1 a2 a3\n" lines and
// clears the log.
MOCK_API size_t mock_cuda_take_launches(char* out, size_t capacity) {
  std::lock_guard<std::mutex> lock(g_mutex);
  std::string text;
  for (const Launch& l : g_launches) {
    char line[384];
    snprintf(line, sizeof(line), "%s %u %u %u %u %u %u %u %llu %llu %llu %llu\n", l.name.c_str(),
             l.grid[0], l.grid[1], l.grid[2], l.block[0], l.block[1], l.block[2], l.shared,
             (unsigned long long)l.args[0], (unsigned long long)l.args[1],
             (unsigned long long)l.args[2], (unsigned long long)l.args[3]);
    text += line;
  }
  g_launches.clear();
  const size_t n = text.size() < capacity ? text.size() : capacity;
  memcpy(out, text.data(), n);
  return n;
}

// The arguments of every cuTensorMapEncodeTiled call since the last take, as text lines
// "type rank swizzle base d0..d4 s0..s4 b0..b4".
MOCK_API size_t mock_cuda_take_tensor_maps(char* out, size_t capacity) {
  std::lock_guard<std::mutex> lock(g_mutex);
  std::string text;
  for (const EncodedMap& m : g_maps) {
    char line[512];
    int at = snprintf(line, sizeof(line), "%u %u %u %llu", m.type, m.rank, m.swizzle,
                      (unsigned long long)m.base);
    for (int i = 0; i < 5; ++i) at += snprintf(line + at, sizeof(line) - at, " %llu", (unsigned long long)m.dims[i]);
    for (int i = 0; i < 5; ++i) at += snprintf(line + at, sizeof(line) - at, " %llu", (unsigned long long)m.strides[i]);
    for (int i = 0; i < 5; ++i) at += snprintf(line + at, sizeof(line) - at, " %u", m.box[i]);
    text += line;
    text += "\n";
  }
  g_maps.clear();
  const size_t n = text.size() < capacity ? text.size() : capacity;
  memcpy(out, text.data(), n);
  return n;
}

// ---- the driver API subset the launch layer resolves (hal/internal.h B200_CUDA_SYMBOLS) -----

MOCK_API CUresult cuInit(unsigned) { return CUDA_SUCCESS; }
MOCK_API CUresult cuDriverGetVersion(int* v) { *v = 12090; return CUDA_SUCCESS; }
MOCK_API CUresult cuGetErrorName(CUresult r, const char** s) {
  *s = r == CUDA_SUCCESS ? "CUDA_SUCCESS" : "CUDA_ERROR_MOCK";
  return CUDA_SUCCESS;
}
MOCK_API CUresult cuGetErrorString(CUresult r, const char** s) {
  *s = r == CUDA_SUCCESS ? "no error" : "mock driver error";
  return CUDA_SUCCESS;
}
MOCK_API CUresult cuDeviceGet(CUdevice* d, int ordinal) {
  if (ordinal != 0) return CUDA_ERROR_INVALID_DEVICE;
  *d = 0;
  return CUDA_SUCCESS;
}
MOCK_API CUresult cuDeviceGetCount(int* n) { *n = 1; return CUDA_SUCCESS; }
MOCK_API CUresult cuDeviceGetName(char* name, int len, CUdevice) {
  snprintf(name, (size_t)len, "mock B200 (no GPU)");
  return CUDA_SUCCESS;
}
MOCK_API CUresult cuDeviceGetAttribute(int* v, CUdevice_attribute a, CUdevice) {
  switch (a) {
    case CU_DEVICE_ATTRIBUTE_COMPUTE_CAPABILITY_MAJOR: *v = 10; break;
    case CU_DEVICE_ATTRIBUTE_COMPUTE_CAPABILITY_MINOR: *v = 0; break;
    case CU_DEVICE_ATTRIBUTE_MULTIPROCESSOR_COUNT: *v = 148; break;
    case CU_DEVICE_ATTRIBUTE_MAX_SHARED_MEMORY_PER_BLOCK_OPTIN: *v = 232448; break;
    case CU_DEVICE_ATTRIBUTE_MAX_BLOCK_DIM_X: *v = 1024; break;
    case CU_DEVICE_ATTRIBUTE_MAX_BLOCK_DIM_Y: *v = 1024; break;
    case CU_DEVICE_ATTRIBUTE_MAX_BLOCK_DIM_Z: *v = 64; break;
    case CU_DEVICE_ATTRIBUTE_MAX_THREADS_PER_BLOCK: *v = 1024; break;
    default: *v = 0; break;
  }
  return CUDA_SUCCESS;
}
MOCK_API CUresult cuDevicePrimaryCtxRetain(CUcontext* c, CUdevice) {
  *c = reinterpret_cast<CUcontext>(&g_dummy_ctx);
  return CUDA_SUCCESS;
}
MOCK_API CUresult cuDevicePrimaryCtxRelease_v2(CUdevice) { return CUDA_SUCCESS; }
MOCK_API CUresult cuCtxSetCurrent(CUcontext) { return CUDA_SUCCESS; }
MOCK_API CUresult cuCtxGetCurrent(CUcontext* c) {
  *c = reinterpret_cast<CUcontext>(&g_dummy_ctx);
  return CUDA_SUCCESS;
}
MOCK_API CUresult cuCtxSynchronize() { return CUDA_SUCCESS; }
MOCK_API CUresult cuStreamCreateWithPriority(CUstream* s, unsigned, int) {
  *s = reinterpret_cast<CUstream>(new int(0));
  return CUDA_SUCCESS;
}
MOCK_API CUresult cuStreamCreate(CUstream* s, unsigned) { return cuStreamCreateWithPriority(s, 0, 0); }
MOCK_API CUresult cuStreamDestroy_v2(CUstream s) {
  delete reinterpret_cast<int*>(s);
  return CUDA_SUCCESS;
}
MOCK_API CUresult cuStreamSynchronize(CUstream) { return CUDA_SUCCESS; }
MOCK_API CUresult cuStreamWaitEvent(CUstream, CUevent, unsigned) { return CUDA_SUCCESS; }
MOCK_API CUresult cuStreamWriteValue32_v2(CUstream, CUdeviceptr p, cuuint32_t v, unsigned) {
  *reinterpret_cast<uint32_t*>(p) = v;
  return CUDA_SUCCESS;
}
MOCK_API CUresult cuCtxGetStreamPriorityRange(int* least, int* greatest) {
  *least = 0;
  *greatest = -5;
  return CUDA_SUCCESS;
}
MOCK_API CUresult cuEventCreate(CUevent* e, unsigned) {
  *e = reinterpret_cast<CUevent>(new int(0));
  return CUDA_SUCCESS;
}
MOCK_API CUresult cuEventDestroy_v2(CUevent e) {
  delete reinterpret_cast<int*>(e);
  return CUDA_SUCCESS;
}
MOCK_API CUresult cuEventRecord(CUevent, CUstream) { return CUDA_SUCCESS; }
MOCK_API CUresult cuEventQuery(CUevent) { return CUDA_SUCCESS; }
MOCK_API CUresult cuEventSynchronize(CUevent) { return CUDA_SUCCESS; }
MOCK_API CUresult cuEventElapsedTime(float* ms, CUevent, CUevent) { *ms = 0.f; return CUDA_SUCCESS; }
MOCK_API CUresult cuEventElapsedTime_v2(float* ms, CUevent, CUevent) { *ms = 0.f; return CUDA_SUCCESS; }
MOCK_API CUresult cuMemAlloc_v2(CUdeviceptr* p, size_t n) {
  void* m = nullptr;
  if (posix_memalign(&m, 512, n ? n : 1) != 0) return CUDA_ERROR_OUT_OF_MEMORY;
  memset(m, 0xCD, n);  // "uninitialised device memory"
  *p = reinterpret_cast<CUdeviceptr>(m);
  return CUDA_SUCCESS;
}
MOCK_API CUresult cuMemFree_v2(CUdeviceptr p) {
  free(reinterpret_cast<void*>(p));
  return CUDA_SUCCESS;
}
MOCK_API CUresult cuMemHostAlloc(void** p, size_t n, unsigned) {
  return posix_memalign(p, 512, n ? n : 1) == 0 ? CUDA_SUCCESS : CUDA_ERROR_OUT_OF_MEMORY;
}
MOCK_API CUresult cuMemFreeHost(void* p) {
  free(p);
  return CUDA_SUCCESS;
}
MOCK_API CUresult cuMemHostGetDevicePointer_v2(CUdeviceptr* d, void* p, unsigned) {
  *d = reinterpret_cast<CUdeviceptr>(p);
  return CUDA_SUCCESS;
}
MOCK_API CUresult cuMemcpyHtoDAsync_v2(CUdeviceptr d, const void* s, size_t n, CUstream) {
  memcpy(reinterpret_cast<void*>(d), s, n);
  return CUDA_SUCCESS;
}
MOCK_API CUresult cuMemcpyDtoHAsync_v2(void* d, CUdeviceptr s, size_t n, CUstream) {
  memcpy(d, reinterpret_cast<const void*>(s), n);
  return CUDA_SUCCESS;
}
MOCK_API CUresult cuMemcpyAsync(CUdeviceptr d, CUdeviceptr s, size_t n, CUstream) {
  memmove(reinterpret_cast<void*>(d), reinterpret_cast<const void*>(s), n);
  return CUDA_SUCCESS;
}
MOCK_API CUresult cuMemcpy2DAsync_v2(const CUDA_MEMCPY2D* p, CUstream) {
  const uint8_t* src = p->srcMemoryType == CU_MEMORYTYPE_HOST
                           ? static_cast<const uint8_t*>(p->srcHost)
                           : reinterpret_cast<const uint8_t*>(p->srcDevice);
  uint8_t* dst = p->dstMemoryType == CU_MEMORYTYPE_HOST ? static_cast<uint8_t*>(p->dstHost)
                                                        : reinterpret_cast<uint8_t*>(p->dstDevice);
  for (size_t r = 0; r < p->Height; ++r) {
    memcpy(dst + (p->dstY + r) * p->dstPitch + p->dstXInBytes,
           src + (p->srcY + r) * p->srcPitch + p->srcXInBytes, p->WidthInBytes);
  }
  return CUDA_SUCCESS;
}
MOCK_API CUresult cuMemsetD8Async(CUdeviceptr d, unsigned char v, size_t n, CUstream) {
  do_memset(d, v, 1, n);
  return CUDA_SUCCESS;
}
MOCK_API CUresult cuMemsetD16Async(CUdeviceptr d, unsigned short v, size_t n, CUstream) {
  do_memset(d, v, 2, n);
  return CUDA_SUCCESS;
}
MOCK_API CUresult cuMemsetD32Async(CUdeviceptr d, unsigned v, size_t n, CUstream) {
  do_memset(d, v, 4, n);
  return CUDA_SUCCESS;
}
MOCK_API CUresult cuModuleLoadDataEx(CUmodule* m, const void* image, unsigned, CUjit_option*, void**) {
  const auto* bytes = static_cast<const uint8_t*>(image);
  if (memcmp(bytes, ELFMAG, SELFMAG) != 0) return CUDA_ERROR_INVALID_IMAGE;
  const auto* eh = reinterpret_cast<const Elf64_Ehdr*>(bytes);
  // The image's extent: section headers sit at the end of a cubin.
  const size_t size = (size_t)eh->e_shoff + (size_t)eh->e_shnum * sizeof(Elf64_Shdr);
  if (eh->e_shoff == 0 || size > (64u << 20)) return CUDA_ERROR_INVALID_IMAGE;
  auto* module = new Module();
  module->image.assign(bytes, bytes + size);
  *m = reinterpret_cast<CUmodule>(module);
  return CUDA_SUCCESS;
}
MOCK_API CUresult cuModuleUnload(CUmodule m) {
  delete reinterpret_cast<Module*>(m);
  return CUDA_SUCCESS;
}
MOCK_API CUresult cuModuleGetFunction(CUfunction* f, CUmodule m, const char* name) {
  uint8_t* at = nullptr;
  size_t size = 0;
  if (!find_symbol(reinterpret_cast<Module*>(m)->image, name, &at, &size)) {
    return CUDA_ERROR_NOT_FOUND;
  }
  *f = reinterpret_cast<CUfunction>(new Function{name});  // lives as long as the process
  return CUDA_SUCCESS;
}
MOCK_API CUresult cuModuleGetGlobal_v2(CUdeviceptr* p, size_t* bytes, CUmodule m, const char* name) {
  uint8_t* at = nullptr;
  size_t size = 0;
  if (!find_symbol(reinterpret_cast<Module*>(m)->image, name, &at, &size)) {
    return CUDA_ERROR_NOT_FOUND;
  }
  *p = reinterpret_cast<CUdeviceptr>(at);
  *bytes = size;
  return CUDA_SUCCESS;
}
MOCK_API CUresult cuFuncSetAttribute(CUfunction, CUfunction_attribute, int) { return CUDA_SUCCESS; }
MOCK_API CUresult cuFuncGetAttribute(int* v, CUfunction_attribute, CUfunction) {
  *v = 0;
  return CUDA_SUCCESS;
}
MOCK_API CUresult cuLaunchKernel(CUfunction f, unsigned gx, unsigned gy, unsigned gz, unsigned bx,
                                 unsigned by, unsigned bz, unsigned shared, CUstream, void** params,
                                 void**) {
  return record_launch(reinterpret_cast<Function*>(f), gx, gy, gz, bx, by, bz, shared, params);
}
MOCK_API CUresult cuTensorMapEncodeTiled(CUtensorMap* map, CUtensorMapDataType type, cuuint32_t rank,
                                         void* base, const cuuint64_t* dims,
                                         const cuuint64_t* strides, const cuuint32_t* box,
                                         const cuuint32_t* elem_strides, CUtensorMapInterleave,
                                         CUtensorMapSwizzle swizzle, CUtensorMapL2promotion,
                                         CUtensorMapFloatOOBfill) {
  if (rank < 1 || rank > 5 || reinterpret_cast<uintptr_t>(base) % 16) return CUDA_ERROR_INVALID_VALUE;
  EncodedMap e = {};
  e.type = (uint32_t)type;
  e.rank = rank;
  e.swizzle = (uint32_t)swizzle;
  e.base = reinterpret_cast<uint64_t>(base);
  for (cuuint32_t i = 0; i < rank; ++i) {
    if (dims[i] == 0 || box[i] == 0 || box[i] > 256 || elem_strides[i] != 1) return CUDA_ERROR_INVALID_VALUE;
    e.dims[i] = dims[i];
    e.box[i] = box[i];
    if (i + 1 < rank) {
      if (strides[i] % 16) return CUDA_ERROR_INVALID_VALUE;
      e.strides[i] = strides[i];
    }
  }
  memset(map, 0, sizeof(*map));
  memcpy(map, &e, sizeof(e) < sizeof(*map) ? sizeof(e) : sizeof(*map));
  std::lock_guard<std::mutex> lock(g_mutex);
  g_maps.push_back(e);
  return CUDA_SUCCESS;
}

// ---- graphs: nodes replay in insertion order ----------------------------------------------------

MOCK_API CUresult cuGraphCreate(CUgraph* g, unsigned) {
  *g = reinterpret_cast<CUgraph>(new Graph());
  return CUDA_SUCCESS;
}
MOCK_API CUresult cuGraphDestroy(CUgraph g) {
  delete reinterpret_cast<Graph*>(g);
  return CUDA_SUCCESS;
}
static CUresult add_node(CUgraphNode* node, CUgraph g, std::function<void()> fn) {
  auto* graph = reinterpret_cast<Graph*>(g);
  graph->nodes.push_back(std::move(fn));
  *node = reinterpret_cast<CUgraphNode>(graph->nodes.size());  // 1-based index as the handle
  return CUDA_SUCCESS;
}
MOCK_API CUresult cuGraphAddKernelNode_v2(CUgraphNode* node, CUgraph g, const CUgraphNode*, size_t,
                                          const CUDA_KERNEL_NODE_PARAMS* p) {
  const CUDA_KERNEL_NODE_PARAMS k = *p;
  return add_node(node, g, [k] {
    // The parameter array lives in the command buffer that owns the graph.
    record_launch(reinterpret_cast<Function*>(k.func), k.gridDimX, k.gridDimY, k.gridDimZ,
                  k.blockDimX, k.blockDimY, k.blockDimZ, k.sharedMemBytes, k.kernelParams);
  });
}
MOCK_API CUresult cuGraphAddMemsetNode(CUgraphNode* node, CUgraph g, const CUgraphNode*, size_t,
                                       const CUDA_MEMSET_NODE_PARAMS* p, CUcontext) {
  const CUDA_MEMSET_NODE_PARAMS m = *p;
  if (m.elementSize != 1 && m.elementSize != 2 && m.elementSize != 4) return CUDA_ERROR_INVALID_VALUE;
  return add_node(node, g, [m] { do_memset(m.dst, m.value, m.elementSize, m.width * m.height); });
}
MOCK_API CUresult cuGraphAddMemcpyNode(CUgraphNode* node, CUgraph g, const CUgraphNode*, size_t,
                                       const CUDA_MEMCPY3D* p, CUcontext) {
  const CUDA_MEMCPY3D c = *p;
  if (c.Height != 1 || c.Depth != 1) return CUDA_ERROR_NOT_SUPPORTED;
  return add_node(node, g, [c] {
    const uint8_t* src = c.srcMemoryType == CU_MEMORYTYPE_HOST
                             ? static_cast<const uint8_t*>(c.srcHost)
                             : reinterpret_cast<const uint8_t*>(c.srcDevice);
    uint8_t* dst = c.dstMemoryType == CU_MEMORYTYPE_HOST ? static_cast<uint8_t*>(c.dstHost)
                                                          : reinterpret_cast<uint8_t*>(c.dstDevice);
    memmove(dst + c.dstXInBytes, src + c.srcXInBytes, c.WidthInBytes);
  });
}
MOCK_API CUresult cuGraphAddEmptyNode(CUgraphNode* node, CUgraph g, const CUgraphNode*, size_t) {
  return add_node(node, g, [] {});
}
MOCK_API CUresult cuGraphAddChildGraphNode(CUgraphNode* node, CUgraph g, const CUgraphNode*, size_t,
                                           CUgraph child) {
  const std::vector<std::function<void()>> nodes = reinterpret_cast<Graph*>(child)->nodes;
  return add_node(node, g, [nodes] {
    for (const auto& fn : nodes) fn();
  });
}
MOCK_API CUresult cuStreamBeginCapture_v2(CUstream, CUstreamCaptureMode) { return CUDA_SUCCESS; }
MOCK_API CUresult cuStreamEndCapture(CUstream, CUgraph* g) { return cuGraphCreate(g, 0); }
MOCK_API CUresult cuGraphInstantiateWithFlags(CUgraphExec* e, CUgraph g, unsigned long long) {
  *e = reinterpret_cast<CUgraphExec>(new Graph(*reinterpret_cast<Graph*>(g)));
  return CUDA_SUCCESS;
}
MOCK_API CUresult cuGraphExecDestroy(CUgraphExec e) {
  delete reinterpret_cast<Graph*>(e);
  return CUDA_SUCCESS;
}
MOCK_API CUresult cuGraphLaunch(CUgraphExec e, CUstream) {
  for (const auto& fn : reinterpret_cast<Graph*>(e)->nodes) fn();
  return CUDA_SUCCESS;
}
