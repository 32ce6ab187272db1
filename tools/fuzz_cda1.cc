// fuzz_cda1.cc — mutation fuzzer for the untrusted-input side of the launch layer: the CDA1
// header + flatbuffer reader behind iree_b200_executable_infer_format (the same
// read_executable_header / parse_executable_def that prepare_executable runs before it touches
// the driver; reference: native_executable.c:127-133 "treat the flatbuffer as untrusted").
// Needs no GPU. Built with -fsanitize=address,undefined by tools/fuzz_cda1.sh, which links the
// launch layer's sources directly; any out-of-bounds read, misaligned access or overflow aborts.
//
//   fuzz_cda1 <iree_b200_kernels.cda1> [iterations] [seed]

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "iree_b200_hal.h"

static uint64_t state = 0x9E3779B97F4A7C15ull;
static uint64_t next() {  // xorshift64*
  state ^= state >> 12;
  state ^= state << 25;
  state ^= state >> 27;
  return state * 0x2545F4914F6CDD1Dull;
}

int main(int argc, char** argv) {
  if (argc < 2) {
    fprintf(stderr, "usage: %s <blob.cda1> [iterations] [seed]\n", argv[0]);
    return 2;
  }
  FILE* f = fopen(argv[1], "rb");
  if (!f) return 2;
  std::vector<uint8_t> blob;
  uint8_t chunk[65536];
  for (size_t n; (n = fread(chunk, 1, sizeof(chunk), f)) > 0;) blob.insert(blob.end(), chunk, chunk + n);
  fclose(f);
  const long iterations = argc > 2 ? atol(argv[2]) : 200000;
  if (argc > 3) state ^= strtoull(argv[3], nullptr, 0);

  // The module images (cubins) are opaque payload: the interesting bytes are the 64-byte
  // header, the root / export tables at the tail and the vtables; keep the corpus small by
  // working on a copy whose mutations concentrate there.
  char format[16];
  size_t inferred = 0;
  if (iree_b200_executable_infer_format(blob.data(), blob.size(), sizeof(format), format,
                                        &inferred) != IREE_B200_STATUS_OK) {
    fprintf(stderr, "seed blob rejected: %s\n", iree_b200_status_message());
    return 1;
  }
  long accepted = 0, rejected = 0;
  std::vector<uint8_t> work;
  for (long it = 0; it < iterations; ++it) {
    work = blob;
    const int kind = (int)(next() % 6);
    const size_t n = work.size();
    // Flatbuffers are built back to front: tables and vtables sit near the start of the
    // buffer region (right after the header), strings/vectors of exports follow.
    const size_t hot = 64 + (size_t)(next() % 65536) % (n - 64);
    switch (kind) {
      case 0:  // flip bits in the header
        work[next() % 64] ^= (uint8_t)(1u << (next() % 8));
        break;
      case 1:  // overwrite a 32-bit word near the tables with an interesting value
      case 2: {
        static const uint32_t values[] = {0, 1, 4, 0x7fffffff, 0x80000000u, 0xffffffffu,
                                          0xfffffffcu, 65535, 65536};
        uint32_t v = values[next() % 9];
        if (kind == 2) v = (uint32_t)next();
        memcpy(&work[hot & ~(size_t)3], &v, 4 <= n - (hot & ~(size_t)3) ? 4 : 0);
        break;
      }
      case 3:  // truncate
        work.resize(next() % n);
        break;
      case 4: {  // splat a run of random bytes
        const size_t len = 1 + next() % 64;
        for (size_t i = 0; i < len && hot + i < n; ++i) work[hot + i] = (uint8_t)next();
        break;
      }
      default: {  // several 16-bit pokes (vtable entries)
        for (int i = 0; i < 4; ++i) {
          const size_t p = (64 + (size_t)(next() % 8192)) % n & ~(size_t)1;
          const uint16_t v = (uint16_t)next();
          if (p + 2 <= n) memcpy(&work[p], &v, 2);
        }
        break;
      }
    }
    // Give the reader an exactly-sized heap allocation so ASAN sees any overrun.
    uint8_t* exact = work.empty() ? nullptr : (uint8_t*)malloc(work.size());
    if (exact) memcpy(exact, work.data(), work.size());
    const iree_b200_status_t s = iree_b200_executable_infer_format(
        exact, work.size(), sizeof(format), format, &inferred);
    free(exact);
    (s == IREE_B200_STATUS_OK ? accepted : rejected)++;
  }
  printf("fuzz_cda1: %ld inputs, %ld rejected with a status, %ld still valid, no crash\n",
         iterations, rejected, accepted);
  return 0;
}
