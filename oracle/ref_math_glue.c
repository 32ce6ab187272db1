// ref_math_glue.c — exports the reference's own narrow-float conversions
// (runtime/src/iree/base/internal/math.h, header-only C, included from where it lies
// under /root/reference) so tests can pin oracle.c's restatement against them.
// Test infrastructure only; built into oracle/_ref/libiree_uk_ref.so by build_ref.sh.

#include <stdint.h>

#include "iree/base/internal/math.h"

#define ORACLE_REF_API __attribute__((visibility("default")))

ORACLE_REF_API float oracle_ref_f8e4m3fn_to_f32(uint8_t b) { return iree_math_f8e4m3fn_to_f32(b); }
ORACLE_REF_API float oracle_ref_f8e5m2_to_f32(uint8_t b) { return iree_math_f8e5m2_to_f32(b); }
ORACLE_REF_API uint8_t oracle_ref_f32_to_f8e4m3fn(float f) { return iree_math_f32_to_f8e4m3fn(f); }
ORACLE_REF_API uint8_t oracle_ref_f32_to_f8e5m2(float f) { return iree_math_f32_to_f8e5m2(f); }
ORACLE_REF_API float oracle_ref_f16_to_f32(uint16_t b) { return iree_math_f16_to_f32(b); }
ORACLE_REF_API float oracle_ref_bf16_to_f32(uint16_t b) { return iree_math_bf16_to_f32(b); }
ORACLE_REF_API uint16_t oracle_ref_f32_to_f16(float f) { return iree_math_f32_to_f16(f); }
ORACLE_REF_API uint16_t oracle_ref_f32_to_bf16(float f) { return iree_math_f32_to_bf16(f); }

// OCP microscaling element / scale types (math.h:606-643).
ORACLE_REF_API float oracle_ref_f4e2m1fn_to_f32(uint8_t b) { return iree_math_f4e2m1fn_to_f32(b); }
ORACLE_REF_API float oracle_ref_f8e8m0fnu_to_f32(uint8_t b) { return iree_math_f8e8m0fnu_to_f32(b); }
ORACLE_REF_API uint8_t oracle_ref_f32_to_f8e8m0fnu(float f) { return iree_math_f32_to_f8e8m0fnu(f); }

// Bulk forms: |n| values at once (the tests sweep millions of bit patterns).
ORACLE_REF_API void oracle_ref_encode(int kind, int64_t n, const float* in, uint16_t* out) {
  for (int64_t i = 0; i < n; ++i) {
    switch (kind) {
      case 0: out[i] = iree_math_f32_to_f8e4m3fn(in[i]); break;
      case 1: out[i] = iree_math_f32_to_f8e5m2(in[i]); break;
      case 2: out[i] = iree_math_f32_to_f16(in[i]); break;
      case 4: out[i] = iree_math_f32_to_f8e8m0fnu(in[i]); break;
      default: out[i] = iree_math_f32_to_bf16(in[i]); break;
    }
  }
}
