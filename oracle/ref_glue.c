// ref_glue.c — drives the REFERENCE's CPU micro-kernels as a row-major matmul.
//
// Test / baseline infrastructure only (like everything under oracle/): linked with the
// reference's ukernel sources by oracle/build_ref.sh into oracle/_ref/libiree_uk_ref.so.
// The arithmetic is entirely the reference's (iree_uk_query_tile_sizes_2d, iree_uk_pack,
// iree_uk_mmt4d, iree_uk_unpack — what an llvm-cpu module calls when data tiling + ukernels
// are on); this file only sequences them the way
// runtime/src/iree/builtins/ukernel/tools/e2e_matmul_benchmark.c:200-330 does and spreads
// the outer M tiles over OpenMP threads (the llvm-cpu path distributes the same tiles over
// its task workers).

#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "iree/builtins/ukernel/api.h"
#include "iree/schemas/cpu_data.h"

#define ORACLE_REF_API __attribute__((visibility("default")))

static void fill_cpu_data(uint64_t* cpu_data) {
  memset(cpu_data, 0, sizeof(uint64_t) * IREE_CPU_DATA_FIELD_COUNT);
  uint64_t f = 0;
  __builtin_cpu_init();
#define BIT(name, flag) \
  if (__builtin_cpu_supports(name)) f |= (flag)
  BIT("avx", IREE_CPU_DATA0_X86_64_AVX);
  BIT("fma", IREE_CPU_DATA0_X86_64_FMA);
  BIT("f16c", IREE_CPU_DATA0_X86_64_F16C);
  BIT("avx2", IREE_CPU_DATA0_X86_64_AVX2);
  BIT("avx512f", IREE_CPU_DATA0_X86_64_AVX512F);
  BIT("avx512cd", IREE_CPU_DATA0_X86_64_AVX512CD);
  BIT("avx512vl", IREE_CPU_DATA0_X86_64_AVX512VL);
  BIT("avx512dq", IREE_CPU_DATA0_X86_64_AVX512DQ);
  BIT("avx512bw", IREE_CPU_DATA0_X86_64_AVX512BW);
  BIT("avx512vnni", IREE_CPU_DATA0_X86_64_AVX512VNNI);
  BIT("avx512bf16", IREE_CPU_DATA0_X86_64_AVX512BF16);
#undef BIT
  cpu_data[0] = f;
}

ORACLE_REF_API uint64_t oracle_ref_cpu_features(void) {
  uint64_t cpu_data[IREE_CPU_DATA_FIELD_COUNT];
  fill_cpu_data(cpu_data);
  return cpu_data[0];
}

static void query(uint32_t op, uint32_t role, int64_t s0, int64_t s1, const uint64_t* cpu_data,
                  int* t0, int* t1) {
  iree_uk_query_tile_sizes_2d_params_t p = {
      .flags = op | role, .size0 = s0, .size1 = s1, .cpu_data = cpu_data};
  iree_uk_query_tile_sizes_2d_out_params_t o = {0, 0};
  iree_uk_query_tile_sizes_2d(&p, &o);
  *t0 = (int)o.tile_size0;
  *t1 = (int)o.tile_size1;
}

// type: 0 = i8 x i8 -> i32, 1 = f16 x f16 -> f32, 2 = bf16 x bf16 -> f32 (oracle.c's ids).
// lhs [M,K], rhs [K,N], out [M,N] row-major. |force_generic| masks the cpu features so the
// generic C tiles run. Returns 0, or -1 on bad arguments / allocation failure; the chosen
// tile sizes are written to tiles[3] = {M0, N0, K0} when non-null.
ORACLE_REF_API int oracle_ref_matmul(int type, int64_t M, int64_t N, int64_t K, const void* lhs,
                                     const void* rhs, void* out, int force_generic, int* tiles) {
  uint32_t mm, qts, pk, un;
  int in_size;
  switch (type) {
    case 0:
      mm = IREE_UK_FLAG_MMT4D_TYPE_S8S8S32;
      qts = IREE_UK_FLAG_QUERY_TILE_SIZES_OPERATION_MATMUL_I8I8I32;
      pk = IREE_UK_FLAG_PACK_TYPE_I8I8;
      un = IREE_UK_FLAG_UNPACK_TYPE_I32I32;
      in_size = 1;
      break;
    case 1:
      mm = IREE_UK_FLAG_MMT4D_TYPE_F16F16F32;
      qts = IREE_UK_FLAG_QUERY_TILE_SIZES_OPERATION_MATMUL_F16F16F32;
      pk = IREE_UK_FLAG_PACK_TYPE_F16F16;
      un = IREE_UK_FLAG_UNPACK_TYPE_F32F32;
      in_size = 2;
      break;
    case 2:
      mm = IREE_UK_FLAG_MMT4D_TYPE_BF16BF16F32;
      qts = IREE_UK_FLAG_QUERY_TILE_SIZES_OPERATION_MATMUL_BF16BF16F32;
      pk = IREE_UK_FLAG_PACK_TYPE_BF16BF16;
      un = IREE_UK_FLAG_UNPACK_TYPE_F32F32;
      in_size = 2;
      break;
    default:
      return -1;
  }
  if (M <= 0 || N <= 0 || K <= 0) return -1;
  uint64_t cpu_data[IREE_CPU_DATA_FIELD_COUNT];
  fill_cpu_data(cpu_data);
  if (force_generic) cpu_data[0] = 0;

  int M0, K0, N0, k0b, m0b, n0b;
  query(qts, IREE_UK_FLAG_QUERY_TILE_SIZES_OPERAND_ROLE_LHS, M, K, cpu_data, &M0, &K0);
  query(qts, IREE_UK_FLAG_QUERY_TILE_SIZES_OPERAND_ROLE_RHS, K, N, cpu_data, &n0b, &k0b);
  query(qts, IREE_UK_FLAG_QUERY_TILE_SIZES_OPERAND_ROLE_RESULT, M, N, cpu_data, &m0b, &N0);
  if (M0 <= 0 || K0 <= 0 || N0 <= 0 || k0b != K0 || m0b != M0 || n0b != N0) return -1;
  // iree_uk_query_tile_sizes only knows f32 and i8 (query_tile_sizes_x86_64_entry_point.c);
  // for f16/bf16 the tile shape is chosen by the compiler's CPU encoding materialisation to
  // match the arch tiles of arch/x86_64/mmt4d_x86_64_tiles.inl. Restate that choice here.
  {
    const uint64_t f = cpu_data[0];
    const uint64_t avx512_base = IREE_CPU_DATA0_X86_64_AVX512F | IREE_CPU_DATA0_X86_64_AVX512BW |
                                 IREE_CPU_DATA0_X86_64_AVX512DQ | IREE_CPU_DATA0_X86_64_AVX512VL |
                                 IREE_CPU_DATA0_X86_64_AVX512CD;
    const uint64_t avx2_fma = IREE_CPU_DATA0_X86_64_AVX2 | IREE_CPU_DATA0_X86_64_FMA |
                              IREE_CPU_DATA0_X86_64_F16C;
    if (type == 2 && (f & avx512_base) == avx512_base && (f & IREE_CPU_DATA0_X86_64_AVX512BF16)) {
      M0 = 16, N0 = 16, K0 = 2;
    } else if (type == 1 && (f & avx512_base) == avx512_base) {
      M0 = 16, N0 = 16, K0 = 1;
    } else if (type == 1 && (f & avx2_fma) == avx2_fma) {
      M0 = 8, N0 = 8, K0 = 1;
    }
  }
  if (tiles) {
    tiles[0] = M0;
    tiles[1] = N0;
    tiles[2] = K0;
  }
  const int64_t M1 = (M + M0 - 1) / M0, K1 = (K + K0 - 1) / K0, N1 = (N + N0 - 1) / N0;
  const int64_t lhs_stride0 = K1 * M0 * K0, rhs_stride0 = K1 * N0 * K0, out_stride0 = N1 * M0 * N0;
  void* plhs = malloc((size_t)(M1 * lhs_stride0 * in_size));
  void* prhs = malloc((size_t)(N1 * rhs_stride0 * in_size));
  void* pout = malloc((size_t)(M1 * out_stride0 * 4));
  if (!plhs || !prhs || !pout) {
    free(plhs);
    free(prhs);
    free(pout);
    return -1;
  }
  // RHS [K,N] -> [N1,K1,N0,K0] (transposing pack). In a compiled module the weights'
  // packing is hoisted/const-evaluated; it is done (and timed by the caller) here.
#pragma omp parallel for schedule(static)
  for (int64_t n1 = 0; n1 < N1; ++n1) {
    const int64_t n_begin = n1 * N0;
    const int64_t cols = (N - n_begin < N0) ? (N - n_begin) : N0;
    iree_uk_pack(rhs, n_begin, N, 1, prhs, n1 * rhs_stride0, rhs_stride0, (int64_t)N0 * K0, K, cols,
                 1, K1, N0, K0, 0,
                 pk | IREE_UK_FLAG_PACK_TRANSPOSE_INNER | IREE_UK_FLAG_PACK_TRANSPOSE_OUTER,
                 cpu_data);
  }
  // LHS pack, mmt4d and unpack per block of outer M tiles.
  const int64_t block = 8;
#pragma omp parallel for schedule(dynamic, 1)
  for (int64_t m1 = 0; m1 < M1; m1 += block) {
    const int64_t cnt = (M1 - m1 < block) ? (M1 - m1) : block;
    const int64_t row_begin = m1 * M0;
    const int64_t rows = (M - row_begin < cnt * M0) ? (M - row_begin) : cnt * M0;
    iree_uk_pack(lhs, row_begin * K, K, 1, plhs, m1 * lhs_stride0, lhs_stride0, (int64_t)M0 * K0,
                 rows, K, cnt, K1, M0, K0, 0, pk, cpu_data);
    iree_uk_mmt4d(plhs, m1 * lhs_stride0, lhs_stride0, prhs, 0, rhs_stride0, pout,
                  m1 * out_stride0, out_stride0, cnt, N1, K1, M0, N0, K0,
                  mm | IREE_UK_FLAG_MMT4D_ALLOW_GENERIC_FALLBACK_TILE_FUNCTION, cpu_data);
    iree_uk_unpack(pout, m1 * out_stride0, out_stride0, (int64_t)M0 * N0, out, row_begin * N, N, 1,
                   cnt, N1, M0, N0, rows, N, un, cpu_data);
  }
  free(plhs);
  free(prhs);
  free(pout);
  return 0;
}

// linalg.mmt4d straight through the reference micro-kernel on already-packed operands:
// lhs [M1,K1,M0,K0], rhs [N1,K1,N0,K0], out [M1,N1,M0,N0] (out is also the accumulator input
// when |accumulate|). |type| as in oracle_ref_matmul. Returns 0 or -1.
ORACLE_REF_API int oracle_ref_mmt4d(int type, int64_t M1, int64_t N1, int64_t K1, int M0, int N0,
                                    int K0, const void* lhs, const void* rhs, void* out,
                                    int accumulate, int force_generic) {
  uint32_t mm;
  switch (type) {
    case 0: mm = IREE_UK_FLAG_MMT4D_TYPE_S8S8S32; break;
    case 1: mm = IREE_UK_FLAG_MMT4D_TYPE_F16F16F32; break;
    case 2: mm = IREE_UK_FLAG_MMT4D_TYPE_BF16BF16F32; break;
    case 3: mm = IREE_UK_FLAG_MMT4D_TYPE_F32F32F32; break;
    default: return -1;
  }
  if (M1 < 0 || N1 < 0 || K1 < 0 || M0 <= 0 || N0 <= 0 || K0 <= 0) return -1;
  uint64_t cpu_data[IREE_CPU_DATA_FIELD_COUNT];
  fill_cpu_data(cpu_data);
  if (force_generic) cpu_data[0] = 0;
  mm |= IREE_UK_FLAG_MMT4D_ALLOW_GENERIC_FALLBACK_TILE_FUNCTION;
  if (accumulate) mm |= IREE_UK_FLAG_MMT4D_ACCUMULATE;
  iree_uk_mmt4d(lhs, 0, K1 * M0 * K0, rhs, 0, K1 * N0 * K0, out, 0, N1 * M0 * N0, M1, N1, K1, M0, N0,
                K0, mm, cpu_data);
  return 0;
}
