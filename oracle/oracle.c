// oracle.c — CPU restatement of the reference's arithmetic for the dense-contraction
// dispatch path (matmul / batch_matmul / attention).
//
// THIS IS TEST INFRASTRUCTURE. Only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may load it. The product path (iree_b200/) never
// links, imports or calls anything in this directory and has no CPU fallback.
//
// Parity status:
//   * matmul / batch_matmul: PINNED — the reference's own test contract is restated here
//     (input generator + per-element loops) and the reference requires bit-exact results on
//     those inputs (tools/testing/e2e/test_utils.c:26-27,258-290).
//   * attention: PINNED on the 17 golden tensors of tests/e2e/linalg_ext_ops/attention*.mlir
//     (atol 1e-4, CheckOps.td:163-164), extracted by tools/extract_attention_golden.py into
//     tests/golden/attention_golden.json. For the large random Llama shapes there is no
//     reference artefact (iree-e2e-attention-test.cc:336-351 never compares), so the pin
//     there is this restatement + the stated tolerance.
//
//   * MXFP4 matmul (f4E2M1FN x f8E8M0FNU block scales): PINNED — element and scale decodes
//     and the scale encoder against the reference's own math.h (oracle/_ref), the matmul by
//     exactness on the reference generator's inputs. Oracle only so far: no CUDA path yet.
//
// Every function cites the reference file:line (relative to /root/reference) it follows.
// No reference source is copied; the loops are re-derived from the semantics.
//
// Build: see oracle/Makefile (gcc -O3 -fopenmp -ffp-contract=off -shared -fPIC).

#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORACLE_API __attribute__((visibility("default")))

// Element type tags shared with oracle/__init__.py.
enum {
  ORACLE_I8 = 0,
  ORACLE_F16 = 1,
  ORACLE_BF16 = 2,
  ORACLE_F32 = 3,
  ORACLE_I32 = 4,
  ORACLE_F8E4M3FN = 5,
  ORACLE_F8E5M2 = 6,
  ORACLE_U8 = 7,         // random bytes: packed sub-byte MX operands (test_utils.c:376-382)
  ORACLE_F8E8M0FNU = 8,  // MX block scales
};

//===----------------------------------------------------------------------===//
// Float helpers. Semantics of runtime/src/iree/base/internal/math.h:364-460
// (f32 -> narrow float, round-to-nearest-even) and :300-362 (narrow -> f32).
//===----------------------------------------------------------------------===//

static inline uint32_t f32_bits(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  return u;
}
static inline float bits_f32(uint32_t u) {
  float f;
  memcpy(&f, &u, 4);
  return f;
}

ORACLE_API float oracle_bf16_to_f32(uint16_t h) {
  return bits_f32((uint32_t)h << 16);
}

ORACLE_API uint16_t oracle_f32_to_bf16(float f) {
  uint32_t u = f32_bits(f);
  if ((u & 0x7f800000u) == 0x7f800000u) {  // inf / nan
    if (u & 0x007fffffu) return (uint16_t)((u >> 16) | 0x0040u);
    return (uint16_t)(u >> 16);
  }
  uint32_t lsb = (u >> 16) & 1u;
  u += 0x7fffu + lsb;  // RNE
  return (uint16_t)(u >> 16);
}

ORACLE_API float oracle_f16_to_f32(uint16_t h) {
  uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
  uint32_t exp = (h >> 10) & 0x1fu;
  uint32_t man = h & 0x3ffu;
  if (exp == 0) {
    if (man == 0) return bits_f32(sign);
    // subnormal: man * 2^-24
    float v = ldexpf((float)man, -24);
    return (sign ? -v : v);
  }
  if (exp == 31) return bits_f32(sign | 0x7f800000u | (man << 13));
  return bits_f32(sign | ((exp + 112u) << 23) | (man << 13));
}

ORACLE_API uint16_t oracle_f32_to_f16(float f) {
  uint32_t u = f32_bits(f);
  uint16_t sign = (uint16_t)((u >> 16) & 0x8000u);
  uint32_t exp = (u >> 23) & 0xffu;
  uint32_t man = u & 0x7fffffu;
  if (exp == 0xff) return (uint16_t)(sign | 0x7c00u | (man ? 0x200u : 0));
  int e = (int)exp - 127;
  if (e > 15) return (uint16_t)(sign | 0x7c00u);  // overflow -> inf
  if (e >= -14) {
    // normal: keep 10 mantissa bits, RNE on the 13 dropped bits.
    uint32_t m = man;
    uint32_t lsb = (m >> 13) & 1u;
    m += 0xfffu + lsb;
    uint32_t he = (uint32_t)(e + 15);
    if (m & 0x800000u) {  // mantissa overflowed into the exponent
      m = 0;
      he += 1;
      if (he >= 31) return (uint16_t)(sign | 0x7c00u);
    }
    return (uint16_t)(sign | (he << 10) | (m >> 13));
  }
  if (e < -25) return sign;  // underflow to zero
  // subnormal: value = 1.man * 2^e -> mantissa units of 2^-24.
  uint32_t full = man | 0x800000u;
  int shift = 13 + (-14 - e);  // >= 14
  uint32_t lsb = (full >> shift) & 1u;
  uint32_t half = (1u << (shift - 1)) - 1u;
  uint32_t r = (full + half + lsb) >> shift;
  return (uint16_t)(sign | r);
}

// OCP 8-bit floats. Decoding: math.h:300-362 with the parameters of :550-560 — f8E5M2 is
// IEEE-like (5 exponent bits, bias 15, infinities; it is the top byte of an f16);
// f8E4M3FN has bias 7, no infinities, and S.1111.111 as its only NaN. Encoding: the
// round-to-nearest-even truncation of :376-516 — overflow gives Inf for f8E5M2 and NaN
// for f8E4M3FN, f32 subnormals flush to zero, results may be subnormal.
ORACLE_API float oracle_f8e5m2_to_f32(uint8_t b) {
  return oracle_f16_to_f32((uint16_t)((uint16_t)b << 8));
}

ORACLE_API float oracle_f8e4m3fn_to_f32(uint8_t b) {
  uint32_t sign = (uint32_t)(b & 0x80u) << 24;
  uint32_t exp = (b >> 3) & 0xfu, man = b & 0x7u;
  if (exp == 0) {
    float v = ldexpf((float)man, -9);
    return sign ? -v : v;
  }
  if (exp == 15 && man == 7) return bits_f32(sign | 0x7fc00000u);
  return bits_f32(sign | ((exp + 120u) << 23) | (man << 20));
}

// Shared RNE encoder for a format with |eb| exponent and |mb| mantissa bits.
static uint8_t f32_to_f8(float f, int eb, int mb, int have_inf) {
  uint32_t u = f32_bits(f);
  uint8_t sign = (uint8_t)((u >> 24) & 0x80u);
  uint32_t fexp = (u >> 23) & 0xffu, fman = u & 0x7fffffu;
  const int bias = (1 << (eb - 1)) - 1;
  const uint8_t exp_mask = (uint8_t)(((1u << eb) - 1u) << mb);
  const uint8_t man_mask = (uint8_t)((1u << mb) - 1u);
  const uint8_t nan = (uint8_t)(sign | exp_mask | man_mask);
  const uint8_t inf = have_inf ? (uint8_t)(sign | exp_mask) : nan;
  if (fexp == 0xff) return fman ? nan : inf;
  if (fexp == 0) return sign;  // zero / f32 subnormal
  int e = (int)fexp - 127;
  if (e > (1 << (eb - 1)) - have_inf) return inf;
  uint32_t full = fman | 0x800000u;
  if (e + bias <= 0) {
    // Subnormal result: units of 2^(1 - bias - mb).
    int shift = (23 - mb) + (1 - bias - e);
    if (shift > 24) return sign;
    uint32_t lsb = (full >> shift) & 1u;
    uint32_t r = (full + ((1u << (shift - 1)) - 1u) + lsb) >> shift;
    return (uint8_t)(sign | r);  // r may carry into the smallest normal: same encoding
  }
  int shift = 23 - mb;
  uint32_t lsb = (fman >> shift) & 1u;
  uint32_t m = fman + ((1u << (shift - 1)) - 1u) + lsb;
  if (m & 0x800000u) {
    m = 0;
    ++e;
  }
  uint32_t de = (uint32_t)(e + bias);
  if (de > ((1u << eb) - 1u)) return have_inf ? inf : nan;
  if (have_inf && de == ((1u << eb) - 1u)) return inf;
  return (uint8_t)(sign | (de << mb) | (m >> shift));
}

ORACLE_API uint8_t oracle_f32_to_f8e5m2(float f) { return f32_to_f8(f, 5, 2, 1); }
ORACLE_API uint8_t oracle_f32_to_f8e4m3fn(float f) { return f32_to_f8(f, 4, 3, 0); }

// OCP microscaling (MX) element and scale types.
// f4E2M1FN (math.h:606-613: 2 exponent bits, 1 mantissa bit, bias 1, no Inf / NaN): the 16
// codes are +-{0, 0.5, 1, 1.5, 2, 3, 4, 6}. |code| is the low nibble.
ORACLE_API float oracle_f4e2m1fn_to_f32(uint8_t code) {
  const uint32_t exp = (code >> 1) & 3u, man = code & 1u;
  const float mag = exp == 0 ? 0.5f * (float)man : ldexpf(1.0f + 0.5f * (float)man, (int)exp - 1);
  return (code & 8u) ? -mag : mag;
}

// f8E8M0FNU (math.h:615-624): a bare biased exponent, 2^(b - 127); 0xFF is NaN; there is
// no zero and no sign.
ORACLE_API float oracle_f8e8m0fnu_to_f32(uint8_t b) {
  return b == 0xFF ? bits_f32(0x7fc00000u) : ldexpf(1.0f, (int)b - 127);
}

// math.h:625-643: non-finite -> 0xFF, non-positive -> 0, otherwise the nearest power of two
// with ties (x = 1.5 * 2^e) going up; clamped to [0, 0xFF].
ORACLE_API uint8_t oracle_f32_to_f8e8m0fnu(float v) {
  if (!isfinite(v)) return 0xFF;
  if (v <= 0.f) return 0;
  int e = 0;
  const float m = 2.0f * frexpf(v, &e);  // v = m * 2^(e - 1), m in [1, 2)
  const int biased = (e - 1) + (m >= 1.5f ? 1 : 0) + 127;
  return (uint8_t)(biased < 0 ? 0 : biased > 0xFF ? 0xFF : biased);
}

// Bulk encoder, same kind numbering as oracle/ref_math_glue.c (tests sweep bit patterns).
ORACLE_API void oracle_encode(int kind, int64_t n, const float* in, uint16_t* out) {
  for (int64_t i = 0; i < n; ++i) {
    switch (kind) {
      case 0: out[i] = oracle_f32_to_f8e4m3fn(in[i]); break;
      case 1: out[i] = oracle_f32_to_f8e5m2(in[i]); break;
      case 2: out[i] = oracle_f32_to_f16(in[i]); break;
      case 4: out[i] = oracle_f32_to_f8e8m0fnu(in[i]); break;
      default: out[i] = oracle_f32_to_bf16(in[i]); break;
    }
  }
}

static inline float load_elem(int type, const void* p, int64_t i) {
  switch (type) {
    case ORACLE_F16:
      return oracle_f16_to_f32(((const uint16_t*)p)[i]);
    case ORACLE_BF16:
      return oracle_bf16_to_f32(((const uint16_t*)p)[i]);
    case ORACLE_F32:
      return ((const float*)p)[i];
    case ORACLE_F8E4M3FN:
      return oracle_f8e4m3fn_to_f32(((const uint8_t*)p)[i]);
    case ORACLE_F8E5M2:
      return oracle_f8e5m2_to_f32(((const uint8_t*)p)[i]);
    default:
      return 0.f;
  }
}

// Rounds an f32 value to the given element type and back (the effect of an
// arith.truncf followed by arith.extf). f16*f16 and bf16*bf16 products are exact in
// f32, so "multiply in f32 then round once" equals a correctly rounded narrow multiply.
static inline float round_to_type(int type, float v) {
  switch (type) {
    case ORACLE_F16:
      return oracle_f16_to_f32(oracle_f32_to_f16(v));
    case ORACLE_BF16:
      return oracle_bf16_to_f32(oracle_f32_to_bf16(v));
    default:
      return v;
  }
}

static inline void store_elem(int type, void* p, int64_t i, float v) {
  switch (type) {
    case ORACLE_F16:
      ((uint16_t*)p)[i] = oracle_f32_to_f16(v);
      break;
    case ORACLE_BF16:
      ((uint16_t*)p)[i] = oracle_f32_to_bf16(v);
      break;
    case ORACLE_F32:
      ((float*)p)[i] = v;
      break;
    case ORACLE_F8E4M3FN:
      ((uint8_t*)p)[i] = oracle_f32_to_f8e4m3fn(v);
      break;
    case ORACLE_F8E5M2:
      ((uint8_t*)p)[i] = oracle_f32_to_f8e5m2(v);
      break;
    default:
      break;
  }
}

//===----------------------------------------------------------------------===//
// Input generator. tools/testing/e2e/test_utils.c:359-366 (minstd LCG evaluated in
// uint32 arithmetic: the product wraps mod 2^32 BEFORE the modulus — this is not
// std::minstd_rand), test_utils.h:99-100 (constants), test_utils.c:368-420 (ranges),
// iree-e2e-matmul-test.cc:719-736 (value = state % range + min, one draw per element,
// row-major order), test_utils.c:306-356 (element write = cast of the small integer).
//===----------------------------------------------------------------------===//

static inline uint32_t prng_next(uint32_t* state) {
  *state = (uint32_t)(*state * 48271u) % 2147483647u;
  return *state;
}

static void min_max_for_type(int type, int32_t* mn, int32_t* mx) {
  switch (type) {
    case ORACLE_I8:
    case ORACLE_BF16:
    case ORACLE_F8E4M3FN:
    case ORACLE_F8E5M2:
      *mn = -2, *mx = 2;
      break;
    case ORACLE_F16:
      *mn = -4, *mx = 4;
      break;
    case ORACLE_U8:  // "random bytes" (test_utils.c:376-382)
      *mn = 0, *mx = 255;
      break;
    case ORACLE_F8E8M0FNU:  // scales 1..8 before rounding to a power of two (:421-424)
      *mn = 1, *mx = 8;
      break;
    default:  // f32 / i32
      *mn = -8, *mx = 8;
      break;
  }
}

ORACLE_API void oracle_generate_random_matrix(int type, int64_t count, int32_t seed,
                                              void* out) {
  int32_t mn, mx;
  min_max_for_type(type, &mn, &mx);
  uint32_t range = (uint32_t)(mx - mn + 1);
  uint32_t state = (uint32_t)seed;
  for (int64_t i = 0; i < count; ++i) {
    int32_t v = (int32_t)(prng_next(&state) % range) + mn;
    switch (type) {
      case ORACLE_I8:
        ((int8_t*)out)[i] = (int8_t)v;
        break;
      case ORACLE_I32:
        ((int32_t*)out)[i] = v;
        break;
      case ORACLE_U8:
        ((uint8_t*)out)[i] = (uint8_t)v;
        break;
      case ORACLE_F8E8M0FNU:  // test_utils.c:345-346
        ((uint8_t*)out)[i] = oracle_f32_to_f8e8m0fnu((float)v);
        break;
      default:
        store_elem(type, out, i, (float)v);
        break;
    }
  }
}

// test_utils.c:56-68 — stride of the sampled result check.
ORACLE_API int oracle_calculate_check_every(int64_t tot_elements, int64_t no_div_of,
                                            int32_t max_elements_to_check) {
  int check_every = 1;
  if (max_elements_to_check) {
    check_every =
        (int)((tot_elements + max_elements_to_check - 1) / max_elements_to_check);
    if (check_every < 1) check_every = 1;
    if (check_every > 1)
      while ((no_div_of % check_every) == 0) ++check_every;
  }
  return check_every;
}

//===----------------------------------------------------------------------===//
// matmul / batch_matmul. iree-e2e-matmul-test.cc:28-49 (i8*i8->i32, sign-extended,
// wrapping i32 add), :73-86 (f16->f32), :108-121 (bf16->f32): per output element,
// acc starts at acc_data[m,n] (or 0), k ascending, acc += ext(lhs)*ext(rhs) as a
// separate multiply and add in the accumulator type (built with -ffp-contract=off).
// transpose_rhs selects rhs[n,k] over rhs[k,n]. Batch: :882-911 — each batch is an
// independent matmul over contiguous slices.
//
// Loop order here is m / k / n (row of accumulators) which keeps every element's own
// k-ascending summation order, so results are bit-identical to the reference loop.
//===----------------------------------------------------------------------===//

ORACLE_API void oracle_matmul_f32acc(int type, int64_t batch, int64_t M, int64_t N,
                                     int64_t K, int transpose_rhs, const void* lhs,
                                     const void* rhs, const float* acc, float* out) {
  // Expand rhs to f32 [K,N] once per batch so the inner loop vectorises.
  for (int64_t b = 0; b < batch; ++b) {
    float* rhs_f32 = (float*)malloc(sizeof(float) * (size_t)K * (size_t)N);
#pragma omp parallel for schedule(static)
    for (int64_t k = 0; k < K; ++k) {
      for (int64_t n = 0; n < N; ++n) {
        int64_t idx = transpose_rhs ? (k + n * K) : (n + k * N);
        rhs_f32[k * N + n] = load_elem(type, rhs, b * K * N + idx);
      }
    }
#pragma omp parallel for schedule(dynamic, 4)
    for (int64_t m = 0; m < M; ++m) {
      float* row = out + (b * M + m) * N;
      if (acc) {
        memcpy(row, acc + (b * M + m) * N, sizeof(float) * (size_t)N);
      } else {
        memset(row, 0, sizeof(float) * (size_t)N);
      }
      for (int64_t k = 0; k < K; ++k) {
        float a = load_elem(type, lhs, (b * M + m) * K + k);
        const float* brow = rhs_f32 + k * N;
        for (int64_t n = 0; n < N; ++n) row[n] += a * brow[n];
      }
    }
    free(rhs_f32);
  }
}

// MXFP4 matmul (tests/e2e/matmul/generate_code_mx.py:153-172): operands are f4E2M1FN packed two
// per byte (element 2i in the low nibble, 2i+1 in the high one: the sub-byte order of
// builtins/ukernel/mmt4d_tile_generic.c:38-45), K is split in blocks of |block| (32) elements
// with one f8E8M0FNU scale per (row, block); both operands are [rows, K] (transpose_rhs
// form). Per element the reference program does, in (block, element) ascending order,
//   l = arith.scaling_extf(lhs, lhs_scale)   = f32(f4) * 2^(scale - 127)
//   r = arith.scaling_extf(rhs, rhs_scale)
//   acc = acc + l * r          (mulf, then addf: two f32 roundings)
// which is also what the runner's checker computes from the scaled f32 operands with the
// f32 reference loop (generate_code_mx.py:285-330, iree-e2e-matmul-test.cc:154-160).
ORACLE_API void oracle_matmul_mxfp4_f32(int64_t M, int64_t N, int64_t K, int64_t block,
                                        const uint8_t* lhs, const uint8_t* lhs_scales,
                                        const uint8_t* rhs, const uint8_t* rhs_scales,
                                        const float* acc, float* out) {
  const int64_t blocks = K / block;
  // Scaled operands once, as the checker materialises them.
  float* l32 = (float*)malloc(sizeof(float) * (size_t)(M * K + 1));
  float* r32 = (float*)malloc(sizeof(float) * (size_t)(N * K + 1));
  for (int pass = 0; pass < 2; ++pass) {
    const int64_t rows = pass ? N : M;
    const uint8_t* bytes = pass ? rhs : lhs;
    const uint8_t* scales = pass ? rhs_scales : lhs_scales;
    float* dst = pass ? r32 : l32;
#pragma omp parallel for schedule(static)
    for (int64_t r = 0; r < rows; ++r) {
      for (int64_t k = 0; k < K; ++k) {
        const uint8_t byte = bytes[(r * K + k) / 2];
        const uint8_t code = (k & 1) ? (uint8_t)(byte >> 4) : (uint8_t)(byte & 0xF);
        dst[r * K + k] =
            oracle_f4e2m1fn_to_f32(code) * oracle_f8e8m0fnu_to_f32(scales[r * blocks + k / block]);
      }
    }
  }
#pragma omp parallel for schedule(dynamic, 4)
  for (int64_t m = 0; m < M; ++m) {
    for (int64_t n = 0; n < N; ++n) {
      float a = acc ? acc[m * N + n] : 0.0f;
      const float* lr = l32 + m * K;
      const float* rr = r32 + n * K;
      for (int64_t k = 0; k < K; ++k) a += lr[k] * rr[k];
      out[m * N + n] = a;
    }
  }
  free(l32);
  free(r32);
}

// linalg.mmt4d on data-tiled operands, as the reference's own test states it one result
// element at a time (runtime/src/iree/builtins/ukernel/tools/mmt4d_test.c:31-45 f16,
// :94-108 bf16, :157-169 s8; the tile loops around them :249-367):
//   lhs [M1,K1,M0,K0], rhs [N1,K1,N0,K0], out [M1,N1,M0,N0];
//   acc = accumulate ? out : 0;  for k1, for k0: acc += lhs[m1,k1,m0,k0] * rhs[n1,k1,n0,k0]
// |type| uses this file's ids (I8 -> i32 result, F16 / BF16 / F32 -> f32 result; f32 as
// mmt4d_tile_generic.c's f32f32f32 loop states it: acc += lhs * rhs, one product and one add
// per step). |acc| NULL = zero
// init. Mirrors iree_uk_mmt4d_early (mmt4d.c:106-114): nothing is read or written for an
// empty result; with K1*K0 == 0 the result is acc (or zeros).
ORACLE_API int oracle_mmt4d(int type, int64_t M1, int64_t N1, int64_t K1, int64_t M0, int64_t N0,
                            int64_t K0, const void* lhs, const void* rhs, const void* acc,
                            void* out) {
  if (type != ORACLE_I8 && type != ORACLE_F16 && type != ORACLE_BF16 && type != ORACLE_F32) return -1;
#pragma omp parallel for collapse(2) schedule(static)
  for (int64_t m1 = 0; m1 < M1; ++m1) {
    for (int64_t n1 = 0; n1 < N1; ++n1) {
      for (int64_t m0 = 0; m0 < M0; ++m0) {
        for (int64_t n0 = 0; n0 < N0; ++n0) {
          const int64_t o = ((m1 * N1 + n1) * M0 + m0) * N0 + n0;
          const int64_t l = (m1 * K1 * M0 + m0) * K0, r = (n1 * K1 * N0 + n0) * K0;
          if (type == ORACLE_I8) {
            // Wrapping i32 accumulate (unsigned arithmetic: no UB on overflow).
            uint32_t a = acc ? (uint32_t)((const int32_t*)acc)[o] : 0u;
            for (int64_t k1 = 0; k1 < K1; ++k1) {
              for (int64_t k0 = 0; k0 < K0; ++k0) {
                const int32_t x = ((const int8_t*)lhs)[l + k1 * M0 * K0 + k0];
                const int32_t y = ((const int8_t*)rhs)[r + k1 * N0 * K0 + k0];
                a += (uint32_t)(x * y);
              }
            }
            ((int32_t*)out)[o] = (int32_t)a;
          } else if (type == ORACLE_F32) {
            float a = acc ? ((const float*)acc)[o] : 0.0f;
            for (int64_t k1 = 0; k1 < K1; ++k1) {
              for (int64_t k0 = 0; k0 < K0; ++k0) {
                a += ((const float*)lhs)[l + k1 * M0 * K0 + k0] *
                     ((const float*)rhs)[r + k1 * N0 * K0 + k0];
              }
            }
            ((float*)out)[o] = a;
          } else {
            float a = acc ? ((const float*)acc)[o] : 0.0f;
            for (int64_t k1 = 0; k1 < K1; ++k1) {
              for (int64_t k0 = 0; k0 < K0; ++k0) {
                const uint16_t x = ((const uint16_t*)lhs)[l + k1 * M0 * K0 + k0];
                const uint16_t y = ((const uint16_t*)rhs)[r + k1 * N0 * K0 + k0];
                const float xf = type == ORACLE_F16 ? oracle_f16_to_f32(x) : oracle_bf16_to_f32(x);
                const float yf = type == ORACLE_F16 ? oracle_f16_to_f32(y) : oracle_bf16_to_f32(y);
                a += xf * yf;
              }
            }
            ((float*)out)[o] = a;
          }
        }
      }
    }
  }
  return 0;
}

ORACLE_API void oracle_matmul_i8_i32(int64_t batch, int64_t M, int64_t N, int64_t K,
                                     int transpose_rhs, const int8_t* lhs,
                                     const int8_t* rhs, const int32_t* acc,
                                     int32_t* out) {
  for (int64_t b = 0; b < batch; ++b) {
    const int8_t* rb = rhs + b * K * N;
    int8_t* rhs_kn = NULL;
    if (transpose_rhs) {
      rhs_kn = (int8_t*)malloc((size_t)K * (size_t)N);
#pragma omp parallel for schedule(static)
      for (int64_t k = 0; k < K; ++k)
        for (int64_t n = 0; n < N; ++n) rhs_kn[k * N + n] = rb[k + n * K];
      rb = rhs_kn;
    }
#pragma omp parallel for schedule(dynamic, 4)
    for (int64_t m = 0; m < M; ++m) {
      int32_t* row = out + (b * M + m) * N;
      if (acc) {
        memcpy(row, acc + (b * M + m) * N, sizeof(int32_t) * (size_t)N);
      } else {
        memset(row, 0, sizeof(int32_t) * (size_t)N);
      }
      for (int64_t k = 0; k < K; ++k) {
        // Wrapping add, as arith.addi: do the sum in uint32.
        int32_t a = (int32_t)lhs[(b * M + m) * K + k];
        const int8_t* brow = rb + k * N;
        for (int64_t n = 0; n < N; ++n)
          row[n] = (int32_t)((uint32_t)row[n] + (uint32_t)(a * (int32_t)brow[n]));
      }
    }
    free(rhs_kn);
  }
}

// Sampled single-element forms (what check_matmul_results evaluates for large
// outputs, iree-e2e-matmul-test.cc:256-284 + test_utils.c:41-68).
ORACLE_API float oracle_matmul_f32acc_element(int type, int64_t M, int64_t N, int64_t K,
                                              int transpose_rhs, const void* lhs,
                                              const void* rhs, int64_t m, int64_t n) {
  (void)M;
  float acc = 0.f;
  for (int64_t k = 0; k < K; ++k) {
    int64_t ridx = transpose_rhs ? (k + n * K) : (n + k * N);
    acc += load_elem(type, lhs, m * K + k) * load_elem(type, rhs, ridx);
  }
  return acc;
}

ORACLE_API int32_t oracle_matmul_i8_i32_element(int64_t M, int64_t N, int64_t K,
                                                int transpose_rhs, const int8_t* lhs,
                                                const int8_t* rhs, int64_t m,
                                                int64_t n) {
  (void)M;
  uint32_t acc = 0;
  for (int64_t k = 0; k < K; ++k) {
    int64_t ridx = transpose_rhs ? (k + n * K) : (n + k * N);
    acc += (uint32_t)((int32_t)lhs[m * K + k] * (int32_t)rhs[ridx]);
  }
  return (int32_t)acc;
}

//===----------------------------------------------------------------------===//
// Attention — the llvm-cpu lowering's arithmetic.
//
//   Q [B,M,K1]  K [B,K2,K1]  V [B,K2,N]  O [B,M,N]   (row-major, indexing maps of
//   tests/e2e/attention/generate_e2e_attention_tests.py:273-300 and
//   tests/e2e/linalg_ext_ops/attention.mlir:14-18)
//
// mask_kind: 0 none, 1 boolean (uint8 0/1, [B,M,K2]), 2 float additive ([B,M,K2] f32).
// region_kind: 0 identity, 1 softcap "tanh(x) * region_param"
//   (attention.mlir:129-135).
//===----------------------------------------------------------------------===//

#define LOG2E_D 1.4426950408889634074

static inline float apply_region(int region_kind, float region_param, float s) {
  if (region_kind == 1) return tanhf(s) * region_param;
  return s;
}

// computeQKAndElementwise, AggregatedOpInterfaceImpl.cpp:323-412:
//  :336-340 scale' = scale * log2e evaluated in the SCALE's type;
//  :355-358 (+ :54-57 cast) Q' = convert(scale', Q type) * Q evaluated in Q's type;
//  :371-375 S = sum_k1 Q'*K accumulated in the S element type (f32);
//  :380     S = region(S);
//  :407-409 (+ :239-291) mask: bool -> +0 / +(-inf); float -> + mask*log2e.
static void compute_scaled_q(int type, int scale_type, float scale, int64_t count,
                             const void* q, float* q_scaled) {
  float log2e_s = round_to_type(scale_type, (float)LOG2E_D);
  float s_in = round_to_type(scale_type, scale);
  float sp = round_to_type(scale_type, s_in * log2e_s);
  float sp_q = round_to_type(type, sp);
  for (int64_t i = 0; i < count; ++i) {
    q_scaled[i] = round_to_type(type, sp_q * load_elem(type, q, i));
  }
}

// OnlineAttentionOp::decomposeOperation, AggregatedOpInterfaceImpl.cpp:533-632, run
// over K2 in tiles of k2_tile (the CPU tiling of KernelDispatch.cpp:2163-2290 changes
// only where the online rescale happens), initialised and finalised as
// convertToOnlineAttention does (TileAttention.cpp:78-96 init acc=0, max=-FLT_MAX
// (useOnlyFiniteValue), sum=0; :136-152 O = cast((1/sum) * acc)).
ORACLE_API void oracle_attention_online(int type, int scale_type, int64_t B, int64_t M,
                                        int64_t K1, int64_t K2, int64_t N,
                                        const void* q, const void* k, const void* v,
                                        float scale, int mask_kind, const void* mask,
                                        int region_kind, float region_param,
                                        int64_t k2_tile, void* out) {
  if (k2_tile <= 0 || k2_tile > K2) k2_tile = K2;
  float* q_scaled = (float*)malloc(sizeof(float) * (size_t)(B * M * K1));
  compute_scaled_q(type, scale_type, scale, B * M * K1, q, q_scaled);

#pragma omp parallel
  {
    float* s = (float*)malloc(sizeof(float) * (size_t)k2_tile);
    float* acc = (float*)malloc(sizeof(float) * (size_t)N);
#pragma omp for collapse(2) schedule(dynamic, 8)
    for (int64_t b = 0; b < B; ++b) {
      for (int64_t m = 0; m < M; ++m) {
        const float* qrow = q_scaled + (b * M + m) * K1;
        float run_max = -3.40282347e+38f;  // -FLT_MAX, TileAttention.cpp:82-84
        float run_sum = 0.f;
        for (int64_t n = 0; n < N; ++n) acc[n] = 0.f;
        for (int64_t t0 = 0; t0 < K2; t0 += k2_tile) {
          int64_t tl = (K2 - t0 < k2_tile) ? (K2 - t0) : k2_tile;
          // S tile.
          for (int64_t j = 0; j < tl; ++j) {
            int64_t kk = t0 + j;
            float a = 0.f;
            for (int64_t d = 0; d < K1; ++d)
              a += qrow[d] * load_elem(type, k, (b * K2 + kk) * K1 + d);
            a = apply_region(region_kind, region_param, a);
            if (mask_kind == 1) {
              uint8_t mv = ((const uint8_t*)mask)[(b * M + m) * K2 + kk];
              a = a + (mv ? 0.f : -INFINITY);
            } else if (mask_kind == 2) {
              float mv = ((const float*)mask)[(b * M + m) * K2 + kk];
              a = a + mv * (float)LOG2E_D;
            }
            s[j] = a;
          }
          // :582 newMax = max(oldMax, rowmax S)
          float new_max = run_max;
          for (int64_t j = 0; j < tl; ++j) new_max = fmaxf(new_max, s[j]);
          // :587-588 norm = exp2(oldMax - newMax); :592-593 normSum = norm*oldSum
          float norm = exp2f(run_max - new_max);
          float new_sum = norm * run_sum;
          // :598 P = exp2(S - newMax); :601 newSum = normSum + rowsum(P) (f32 P)
          for (int64_t j = 0; j < tl; ++j) {
            float p = exp2f(s[j] - new_max);
            new_sum += p;
            // :613-618 P truncated to V's element type before P@V.
            s[j] = round_to_type(type, p);
          }
          // :620-626 acc = norm*acc + P@V, f32.
          for (int64_t n = 0; n < N; ++n) acc[n] = norm * acc[n];
          for (int64_t j = 0; j < tl; ++j) {
            float p = s[j];
            int64_t vbase = (b * K2 + t0 + j) * N;
            for (int64_t n = 0; n < N; ++n) acc[n] += p * load_elem(type, v, vbase + n);
          }
          run_max = new_max;
          run_sum = new_sum;
        }
        float rcp = 1.0f / run_sum;  // TileAttention.cpp:141-146
        for (int64_t n = 0; n < N; ++n)
          store_elem(type, out, (b * M + m) * N + n, rcp * acc[n]);
      }
    }
    free(s);
    free(acc);
  }
  free(q_scaled);
}

// AttentionOp::decomposeOperation, AggregatedOpInterfaceImpl.cpp:418-527 — the
// one-shot form (max over all K2, P = exp2(S-max), P /= sum, truncate P, P@V into
// the OUTPUT element type accumulator). Second opinion for the online form.
ORACLE_API void oracle_attention_oneshot(int type, int scale_type, int64_t B, int64_t M,
                                         int64_t K1, int64_t K2, int64_t N,
                                         const void* q, const void* k, const void* v,
                                         float scale, int mask_kind, const void* mask,
                                         int region_kind, float region_param,
                                         void* out) {
  float* q_scaled = (float*)malloc(sizeof(float) * (size_t)(B * M * K1));
  compute_scaled_q(type, scale_type, scale, B * M * K1, q, q_scaled);
#pragma omp parallel
  {
    float* s = (float*)malloc(sizeof(float) * (size_t)K2);
    float* acc = (float*)malloc(sizeof(float) * (size_t)N);
#pragma omp for collapse(2) schedule(dynamic, 8)
    for (int64_t b = 0; b < B; ++b) {
      for (int64_t m = 0; m < M; ++m) {
        const float* qrow = q_scaled + (b * M + m) * K1;
        float mx = -3.40282347e+38f;
        for (int64_t kk = 0; kk < K2; ++kk) {
          float a = 0.f;
          for (int64_t d = 0; d < K1; ++d)
            a += qrow[d] * load_elem(type, k, (b * K2 + kk) * K1 + d);
          a = apply_region(region_kind, region_param, a);
          if (mask_kind == 1) {
            uint8_t mv = ((const uint8_t*)mask)[(b * M + m) * K2 + kk];
            a = a + (mv ? 0.f : -INFINITY);
          } else if (mask_kind == 2) {
            a = a + ((const float*)mask)[(b * M + m) * K2 + kk] * (float)LOG2E_D;
          }
          s[kk] = a;
          mx = fmaxf(mx, a);
        }
        float sum = 0.f;
        for (int64_t kk = 0; kk < K2; ++kk) {
          s[kk] = exp2f(s[kk] - mx);
          sum += s[kk];
        }
        for (int64_t n = 0; n < N; ++n) acc[n] = 0.f;
        for (int64_t kk = 0; kk < K2; ++kk) {
          float p = round_to_type(type, s[kk] / sum);
          for (int64_t n = 0; n < N; ++n) {
            // accumulator is the output element type (:479-481 accInit uses output's
            // element type): round after each multiply-add for narrow outputs.
            acc[n] = round_to_type(
                type, acc[n] + p * load_elem(type, v, (b * K2 + kk) * N + n));
          }
        }
        for (int64_t n = 0; n < N; ++n)
          store_elem(type, out, (b * M + m) * N + n, acc[n]);
      }
    }
    free(s);
    free(acc);
  }
  free(q_scaled);
}

//===----------------------------------------------------------------------===//
// Collectives (iree_hal_collective_op_t, hal/command_buffer.h:380-462) restated for
// the world-size-2 gloo tests: element-wise sum all-reduce and concatenating all-gather.
//===----------------------------------------------------------------------===//

ORACLE_API void oracle_all_reduce_sum_f32(int ranks, int64_t count, const float* in,
                                          float* out) {
  for (int64_t i = 0; i < count; ++i) {
    float a = 0.f;
    for (int r = 0; r < ranks; ++r) a += in[(int64_t)r * count + i];
    out[i] = a;
  }
}

ORACLE_API int oracle_abi_version(void) { return 1; }
