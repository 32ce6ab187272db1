// iree_b200_ops.h — C ABI of the dense-contraction dispatches themselves.
//
// What a compiled IREE module does for a matched op is: look up the export in the
// executable, evaluate the export's `count` region on the host to get the workgroup
// count, and record hal.command_buffer.dispatch with (constants…, bindings…)
// (compiler/src/iree/compiler/Dialect/HAL/Conversion/StreamToHAL/Patterns.cpp:977-1021).
// The selection guards live in the match-and-replace spec (compiler_plugin/). These entry
// points are that logic as host calls, so a runtime without `iree-compile` (this
// container has none) drives exactly the dispatches the plugin would emit:
//   * iree_b200_ops_select_*   : matcher guards + count region → (export, workgroup count)
//   * iree_b200_ops_record_*   : select + iree_b200_command_buffer_dispatch
//   * iree_b200_ops_*_host     : the `iree-run-module` shaped call — host buffers in,
//                                host buffer out, synchronous (H2D, dispatch, D2H).
// Op semantics are the reference's:
//   matmul / batch_matmul  tools/testing/e2e/iree-e2e-matmul-test.cc:28-121,810-919
//   attention              compiler/src/iree/compiler/Dialect/LinalgExt/IR/AggregatedOpInterfaceImpl.cpp:323-412,533-632
//                          compiler/src/iree/compiler/Dialect/LinalgExt/Transforms/TileAttention.cpp:35-155
// There is no CPU fallback: every entry point fails with a status when no B200 is present.

#ifndef IREE_B200_OPS_H_
#define IREE_B200_OPS_H_

#include "iree_b200_hal.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct iree_b200_ops_t iree_b200_ops_t;

// Element types of the operands (result type is implied: f32 for floats, i32 for i8;
// attention returns the operand type).
enum {
  IREE_B200_ELEMENT_TYPE_I8 = 0,
  IREE_B200_ELEMENT_TYPE_F16 = 1,
  IREE_B200_ELEMENT_TYPE_BF16 = 2,
  IREE_B200_ELEMENT_TYPE_F32 = 3,  // attention only (generic kernel; the reference's goldens)
  // OCP 8-bit floats, matmul only, f32 accumulate and result (tcgen05 kind::f8f6f4;
  // reference loops iree-e2e-matmul-test.cc:123-144, decoding base/internal/math.h:550-560).
  IREE_B200_ELEMENT_TYPE_F8E4M3FN = 4,
  IREE_B200_ELEMENT_TYPE_F8E5M2 = 5,
};

enum {
  IREE_B200_MATMUL_FLAG_ACCUMULATE = 1u << 0,     // out = out + lhs·rhs (linalg outs operand)
  IREE_B200_MATMUL_FLAG_TRANSPOSE_RHS = 1u << 1,  // rhs is [N, K] (generate_code.py:172-174)
  // The dispatch also holds the consumer arith.truncf: out has the operands' element type
  // (f16 / bf16 only, round to nearest even; not combinable with ACCUMULATE).
  IREE_B200_MATMUL_FLAG_TRUNCATE_RESULT = 1u << 2,
  // The dispatch also holds the consumer elementwise op (what IREE's
  // dispatch creation fuses behind a contraction, compiler/src/iree/compiler/DispatchCreation):
  //   x = (out +) lhs·rhs + bias[n]        BIAS: one more binding, bias [N] in the accumulator
  //                                        type (f32 / i32), broadcast along batch and M
  //   x = max(x, 0)                        RELU (needs BIAS): arith.maximumf / arith.maxsi —
  //                                        NaN propagates, -0 becomes +0
  // then TRUNCATE_RESULT if set. i8, f16, bf16 (not the f8 types). Selected through
  // iree_b200_ops_select_matmul with these flags; recorded with ..._record_matmul_bias.
  IREE_B200_MATMUL_FLAG_BIAS = 1u << 3,
  IREE_B200_MATMUL_FLAG_RELU = 1u << 4,
  // Diagnostics: force a kernel family regardless of the shape heuristics.
  IREE_B200_MATMUL_FLAG_FORCE_GENERIC = 1u << 8,
  IREE_B200_MATMUL_FLAG_FORCE_1CTA = 1u << 9,
  IREE_B200_MATMUL_FLAG_FORCE_2CTA = 1u << 10,
  // MXFP4 only: take the expand-to-bf16 + bf16 GEMM route instead of the block-scaled kernel.
  IREE_B200_MATMUL_FLAG_MX_EXPAND = 1u << 11,
  // mmt4d only: take the relayout + row-major GEMM route (needs the workspace) even where the
  // tile-direct tcgen05 export applies (comparison runs, tools/mmt4d_bench.py).
  IREE_B200_MATMUL_FLAG_MMT4D_RELAYOUT = 1u << 12,
  // Select the `*_dtm` exports: the same tcgen05 kernels building their tensor maps on the
  // device (tensormap.replace), i.e. the exports an UNMODIFIED IREE CUDA driver can launch —
  // bindings and push constants only, no `__tma` recipe. Every tcgen05 export has such a twin
  // (all element types, truncating and bias / activation forms; gemm_dtm_kernels.cu).
  IREE_B200_MATMUL_FLAG_DEVICE_MAPS = 1u << 13,
};

// The builtin executable ("CDA1" blob linked into this library: one cubin per kernel
// family, built for sm_100a) and its size; also written to disk by the build as
// iree_b200/_build/iree_b200_kernels.cda1 for use as a #hal.executable.object.
IREE_B200_API const void* iree_b200_ops_builtin_executable_data(size_t* out_size);

// Loads the builtin executable through the executable cache (PTXE path).
IREE_B200_API iree_b200_status_t iree_b200_ops_create(iree_b200_device_t* device,
                                                      iree_b200_ops_t** out_ops);
// Same, from a caller-provided CDA1 blob (e.g. read from the .cda1 file).
IREE_B200_API iree_b200_status_t iree_b200_ops_create_from_data(iree_b200_device_t* device,
                                                                const void* data, size_t size,
                                                                iree_b200_ops_t** out_ops);
// Selection without a GPU: parses |data| (NULL = the builtin executable) for export names and
// layouts and answers iree_b200_ops_select_* / iree_b200_ops_executable() lookups as a device
// with |sm_count| SMs would. What a compiler-side tool evaluating the match-and-replace
// spec's `count` regions needs, and what the CPU test suite uses. Every entry point that would
// record or run a kernel returns FAILED_PRECONDITION for such an object: there is no CPU path.
IREE_B200_API iree_b200_status_t iree_b200_ops_create_offline(const void* data, size_t size,
                                                              int32_t sm_count,
                                                              iree_b200_ops_t** out_ops);
IREE_B200_API void iree_b200_ops_release(iree_b200_ops_t* ops);
IREE_B200_API iree_b200_executable_t* iree_b200_ops_executable(iree_b200_ops_t* ops);

// Caps the grid of the persistent GEMM kernels at (SM count - |reserved_sms|) so that
// concurrently running collective kernels (NCCL on the device's collective stream) find
// free SMs; the persistent kernels otherwise own every SM (1 CTA/SM at 227 KB of shared
// memory) and the collective could only start when they retire. 0 restores the default.
IREE_B200_API iree_b200_status_t iree_b200_ops_set_reserved_sms(iree_b200_ops_t* ops,
                                                                int32_t reserved_sms);

// Tuning of the host-buffer calls (iree_b200_ops_matmul_host / _attention_host). Calls that
// move at least |min_bytes| run as a pipeline: the result is cut into |block_m| x |block_n|
// blocks (rounded to multiples of 256; batch entries are separate blocks), operand panels go
// up on an upload stream in the order that enables blocks soonest, each block's dispatch
// runs on the dispatch stream as soon as its panels have landed, and its result comes back
// on a download stream while later panels are still going up (PCIe is full duplex). Below
// |min_bytes|, or when the operands fail the TMA guards, the call is upload / dispatch /
// download in sequence. Defaults: 16 MiB, 2048 x 2048. Results are identical either way.
IREE_B200_API iree_b200_status_t iree_b200_ops_set_host_pipeline(iree_b200_ops_t* ops,
                                                                 uint64_t min_bytes,
                                                                 int32_t block_m,
                                                                 int32_t block_n);

// Number of dispatches recorded through this object since creation (each is one kernel
// launch when the command buffer executes).
IREE_B200_API uint64_t iree_b200_ops_dispatch_count(const iree_b200_ops_t* ops);

//===----------------------------------------------------------------------===//
// linalg.matmul / linalg.batch_matmul
//===----------------------------------------------------------------------===//

typedef struct iree_b200_matmul_selection_t {
  int32_t export_ordinal;
  uint32_t workgroup_count[3];
  uint32_t constants[5];  // M, N, K, batch, kernel flags
} iree_b200_matmul_selection_t;

// Picks the export and grid for C[batch,M,N] = A[batch,M,K]·B[batch,K,N].
// |alignment| = the largest power of two (≤ 256) dividing all three binding addresses.
// TMA-fed tcgen05 kernels need 16-byte aligned bases and row pitches; everything else
// goes to the generic kernel. Pure host logic (no GPU needed beyond |ops| existing).
IREE_B200_API iree_b200_status_t iree_b200_ops_select_matmul(
    const iree_b200_ops_t* ops, int32_t element_type, int64_t batch, int64_t M, int64_t N,
    int64_t K, uint32_t flags, uint32_t alignment, iree_b200_matmul_selection_t* out_selection);

// Records the dispatch into |cb| (which must be recording). batch == 1 is linalg.matmul.
IREE_B200_API iree_b200_status_t iree_b200_ops_record_matmul(
    iree_b200_ops_t* ops, iree_b200_command_buffer_t* cb, int32_t element_type, int64_t batch,
    int64_t M, int64_t N, int64_t K, uint32_t flags, iree_b200_buffer_ref_t lhs,
    iree_b200_buffer_ref_t rhs, iree_b200_buffer_ref_t out);

// As iree_b200_ops_record_matmul for the FLAG_BIAS dispatches (bindings lhs, rhs, bias, out;
// FLAG_BIAS is implied). |bias| holds N accumulator-type values.
IREE_B200_API iree_b200_status_t iree_b200_ops_record_matmul_bias(
    iree_b200_ops_t* ops, iree_b200_command_buffer_t* cb, int32_t element_type, int64_t batch,
    int64_t M, int64_t N, int64_t K, uint32_t flags, iree_b200_buffer_ref_t lhs,
    iree_b200_buffer_ref_t rhs, iree_b200_buffer_ref_t bias, iree_b200_buffer_ref_t out);

// Host-buffer form of the FLAG_BIAS dispatches: upload, one dispatch, download; above the
// iree_b200_ops_set_host_pipeline threshold the same panel pipeline as
// iree_b200_ops_matmul_host, each block dispatched with its columns of |bias|.
IREE_B200_API iree_b200_status_t iree_b200_ops_matmul_bias_host(
    iree_b200_ops_t* ops, int32_t element_type, int64_t batch, int64_t M, int64_t N, int64_t K,
    uint32_t flags, const void* lhs, const void* rhs, const void* bias, void* out);

// Host-buffer form. |out| holds batch*M*N f32 / i32 values (operand-type values with
// FLAG_TRUNCATE_RESULT); with FLAG_ACCUMULATE it is also the accumulator input.
IREE_B200_API iree_b200_status_t iree_b200_ops_matmul_host(iree_b200_ops_t* ops,
                                                           int32_t element_type, int64_t batch,
                                                           int64_t M, int64_t N, int64_t K,
                                                           uint32_t flags, const void* lhs,
                                                           const void* rhs, void* out);

//===----------------------------------------------------------------------===//
// MXFP4 matmul — PENDING ITS FIRST GPU RUN (kernels/mx_kernels.cu)
//===----------------------------------------------------------------------===//
// out[M,N] f32 = (out +) sum_k f32(lhs[m,k]) * 2^(lhs_scales[m,k/32]-127)
//                            * f32(rhs[n,k]) * 2^(rhs_scales[n,k/32]-127)
// as tests/e2e/matmul/generate_code_mx.py:153-172 states it: lhs [M, K/2] and rhs [N, K/2]
// are f4E2M1FN packed two per byte (low nibble first), the scales are f8E8M0FNU, one per
// 32-element block; K must be a multiple of 32. FLAG_ACCUMULATE is the only flag. The
// kernels live in a second builtin executable that is loaded by the first call: a SIMT form
// (this entry point) and a bf16-expansion + tcgen05 form (the _tensor entry point below); the
// tcgen05 kind::mxf4 block-scaled form is next.
#define IREE_B200_MX_BLOCK_SIZE 32

IREE_B200_API iree_b200_status_t iree_b200_ops_record_matmul_mxfp4(
    iree_b200_ops_t* ops, iree_b200_command_buffer_t* cb, int64_t M, int64_t N, int64_t K,
    uint32_t flags, iree_b200_buffer_ref_t lhs, iree_b200_buffer_ref_t lhs_scales,
    iree_b200_buffer_ref_t rhs, iree_b200_buffer_ref_t rhs_scales, iree_b200_buffer_ref_t out);

// Tensor-core route of the same op: both operands are expanded to bf16 in |workspace| (two
// dispatches of iree_b200_matmul_mxfp4_dequant_bf16 — exact for every f4E2M1FN x f8E8M0FNU
// pair, kernels/mx_kernels.cu), then the bf16 tcgen05 NT GEMM runs on them with the f32
// accumulate of the plain matmul: bit-identical to the reference whenever the partial sums
// are exact (they are on the reference generator's inputs), f32 summation error otherwise.
// Taken when iree_b200_ops_matmul_mxfp4_workspace_size(M, N, K) is non-zero (M, N, K >= 128
// and the result pitch passes the TMA guard) and |workspace| (256-byte aligned) holds that
// many bytes; with a NULL workspace or a shape off the route this records the single SIMT
// dispatch of iree_b200_ops_record_matmul_mxfp4. The host-buffer form below makes the same
// choice with a workspace of its own.
IREE_B200_API uint64_t iree_b200_ops_matmul_mxfp4_workspace_size(int64_t M, int64_t N, int64_t K);
IREE_B200_API iree_b200_status_t iree_b200_ops_record_matmul_mxfp4_tensor(
    iree_b200_ops_t* ops, iree_b200_command_buffer_t* command_buffer, int64_t M, int64_t N,
    int64_t K, uint32_t flags, iree_b200_buffer_ref_t lhs, iree_b200_buffer_ref_t lhs_scales,
    iree_b200_buffer_ref_t rhs, iree_b200_buffer_ref_t rhs_scales,
    iree_b200_buffer_ref_t workspace, iree_b200_buffer_ref_t out);

// Bytes of workspace with which iree_b200_ops_record_matmul_mxfp4_tensor takes the fast form
// of the block-scaled tcgen05 kernel (scales repacked into the instruction's atom order by
// two small dispatches, then moved by TMA); 0 when the shape does not qualify (K % 256,
// N % 4). A workspace of max(this, iree_b200_ops_matmul_mxfp4_workspace_size) serves every
// route.
IREE_B200_API uint64_t iree_b200_ops_matmul_mxfp4_scale_workspace_size(int64_t M, int64_t N,
                                                                       int64_t K);

IREE_B200_API iree_b200_status_t iree_b200_ops_matmul_mxfp4_host(
    iree_b200_ops_t* ops, int64_t M, int64_t N, int64_t K, uint32_t flags, const void* lhs,
    const void* lhs_scales, const void* rhs, const void* rhs_scales, void* out);

//===----------------------------------------------------------------------===//
// linalg.mmt4d — the contraction on data-tiled (pre-packed) operands (SURVEY §8(f) rank 3;
// PENDING its first GPU run — see kernels/mmt4d_kernels.cu)
//===----------------------------------------------------------------------===//
//
// What IREE's encoding path turns a matmul into once the operands carry a tiled layout;
// the CPU backends lower it to the iree_uk_mmt4d micro-kernel
// (runtime/src/iree/builtins/ukernel/mmt4d.h:14-22), whose argument meaning this mirrors:
//   lhs [M1, K1, M0, K0], rhs [N1, K1, N0, K0], out [M1, N1, M0, N0], all dense;
//   out = (FLAG_ACCUMULATE ? out : 0) + sum_{k1,k0} lhs[m1,k1,m0,k0] * rhs[n1,k1,n0,k0]
// in f32 (F16 / BF16 operands) or wrapping i32 (I8). Tile sizes are run-time values in
// [1, 32767] (the micro-kernel's int16 range, mmt4d.c:12-52). No other flag applies. An
// empty result (M1 == 0 or N1 == 0) records nothing; K1 == 0 stores zeros (or leaves out
// unchanged with FLAG_ACCUMULATE), as iree_uk_mmt4d_early does (mmt4d.c:106-114).
IREE_B200_API iree_b200_status_t iree_b200_ops_record_mmt4d(
    iree_b200_ops_t* ops, iree_b200_command_buffer_t* command_buffer, int32_t element_type,
    int64_t M1, int64_t N1, int64_t K1, int32_t M0, int32_t N0, int32_t K0, uint32_t flags,
    iree_b200_buffer_ref_t lhs, iree_b200_buffer_ref_t rhs, iree_b200_buffer_ref_t out);

// Tensor-core route: lhs, rhs (and with FLAG_ACCUMULATE the accumulator) are unpacked to
// row-major in |workspace| (iree_b200_mmt4d_relayout_b{8,16,32}: pure data movement), the
// tcgen05 NT GEMM runs on them, and the result is packed back into [M1, N1, M0, N0] tiles —
// 4 or 5 dispatches. Taken when iree_b200_ops_mmt4d_workspace_size(...) is non-zero (logical
// M, N, K >= 128 and the GEMM's TMA guards hold) and |workspace| (256-byte aligned) holds
// that many bytes; otherwise the single SIMT dispatch of iree_b200_ops_record_mmt4d. Same
// result bits as the plain matmul of the unpacked operands.
IREE_B200_API uint64_t iree_b200_ops_mmt4d_workspace_size(int32_t element_type, int64_t M1,
                                                          int64_t N1, int64_t K1, int32_t M0,
                                                          int32_t N0, int32_t K0);
IREE_B200_API iree_b200_status_t iree_b200_ops_record_mmt4d_tensor(
    iree_b200_ops_t* ops, iree_b200_command_buffer_t* command_buffer, int32_t element_type,
    int64_t M1, int64_t N1, int64_t K1, int32_t M0, int32_t N0, int32_t K0, uint32_t flags,
    iree_b200_buffer_ref_t lhs, iree_b200_buffer_ref_t rhs, iree_b200_buffer_ref_t workspace,
    iree_b200_buffer_ref_t out);

// Host-buffer form (upload, dispatch(es) — tensor-core route when the shape allows —, download).
IREE_B200_API iree_b200_status_t iree_b200_ops_mmt4d_host(
    iree_b200_ops_t* ops, int32_t element_type, int64_t M1, int64_t N1, int64_t K1, int32_t M0,
    int32_t N0, int32_t K0, uint32_t flags, const void* lhs, const void* rhs, void* out);

//===----------------------------------------------------------------------===//
// iree_linalg_ext.attention
//===----------------------------------------------------------------------===//
// Q [B,M,K1], K [B,K2,K1], V [B,K2,N], optional mask [B,M,K2], O [B,M,N]; all operands and
// the result share |element_type| (f16, bf16, or f32 on the generic kernel), the scale is a scalar of the same type
// (tests/e2e/attention/generate_e2e_attention_tests.py:271-300). The region is identity.

enum {
  IREE_B200_ATTENTION_MASK_NONE = 0,
  IREE_B200_ATTENTION_MASK_BOOL = 1,  // i1 stored one byte per element
  IREE_B200_ATTENTION_MASK_F32 = 2,   // additive, natural-log domain (scaled by log2e inside)
  // The causal i1 mask, mask[m, k2] = k2 <= m + (K2 - M) (K2 >= M), WITHOUT a mask operand:
  // what a matcher emits when the op's mask is the iota-compare pattern. The kernels compute
  // it in registers and skip K2 blocks that are masked for a whole tile (tcgen05 prefill:
  // half the work at M == K2). Same result as passing that mask as MASK_BOOL.
  IREE_B200_ATTENTION_MASK_CAUSAL = 3,
};

enum {
  // Diagnostics: force a kernel family regardless of the shape heuristics.
  IREE_B200_ATTENTION_FLAG_FORCE_GENERIC = 1u << 8,
  IREE_B200_ATTENTION_FLAG_FORCE_PREFILL = 1u << 9,
  IREE_B200_ATTENTION_FLAG_FORCE_DECODE = 1u << 10,
  // Within the decode family: the two-dispatch split + merge form / the single-dispatch
  // cluster form (otherwise chosen by the number of (batch, m) rows and K2).
  IREE_B200_ATTENTION_FLAG_FORCE_DECODE_SPLIT = 1u << 11,
  IREE_B200_ATTENTION_FLAG_FORCE_DECODE_CLUSTER = 1u << 12,
  // Prefill kernel: how many of every four exp2 run as a polynomial on the FMA pipe instead
  // of on the MUFU (the prefill kernel's co-limiter): none / 1 of 4 / 2 of 4. Without any of
  // these the default chosen from the measurements applies.
  IREE_B200_ATTENTION_FLAG_EXP2_MUFU_ONLY = 1u << 13,
  IREE_B200_ATTENTION_FLAG_EXP2_POLY_25 = 1u << 14,
  IREE_B200_ATTENTION_FLAG_EXP2_POLY_50 = 1u << 15,
  // Prefill kernel: keep the plain [ceil(M/256), B] grid even when the last wave of equal
  // units would be mostly empty (default: cut that wave's units over K2 and merge).
  IREE_B200_ATTENTION_FLAG_NO_TAIL_SPLIT = 1u << 16,
  // Prefill kernel: the `*_dtm` export, which builds its tensor maps on the device and so is
  // launchable by an unmodified IREE CUDA driver (see IREE_B200_MATMUL_FLAG_DEVICE_MAPS).
  // Unmasked or MASK_CAUSAL; always one dispatch.
  IREE_B200_ATTENTION_FLAG_DEVICE_MAPS = 1u << 17,
};

#define IREE_B200_ATTENTION_MAX_DISPATCHES 2

typedef struct iree_b200_attention_dispatch_t {
  int32_t export_ordinal;
  uint32_t workgroup_count[3];
  uint32_t constant_count;
  uint32_t constants[8];
} iree_b200_attention_dispatch_t;

typedef struct iree_b200_attention_selection_t {
  int32_t dispatch_count;  // 0 (empty result), 1, or 2 (decode split + merge; prefill tail split + merge)
  iree_b200_attention_dispatch_t dispatches[IREE_B200_ATTENTION_MAX_DISPATCHES];
  uint64_t workspace_bytes;  // transient f32 buffer the two-dispatch form needs (else 0)
} iree_b200_attention_selection_t;

// Matcher guards + count regions. tcgen05 prefill kernel: K1 == N == 128, no mask, M > 16,
// 16-byte aligned bindings. Decode kernels: K1 == N == 128, no mask, M <= 16 — one dispatch
// (a cluster of 8 CTAs per (batch, m) row, merged over distributed shared memory) when
// there are enough rows to fill the GPU that way, else split-KV + merge (two dispatches).
// Everything else within K1, N <= 512 goes to the generic kernel; beyond that the status is
// UNIMPLEMENTED (the compiler plugin leaves such ops to IREE's own codegen).
IREE_B200_API iree_b200_status_t iree_b200_ops_select_attention(
    const iree_b200_ops_t* ops, int32_t element_type, int64_t B, int64_t M, int64_t K1,
    int64_t K2, int64_t N, float scale, int32_t mask_kind, uint32_t flags, uint32_t alignment,
    iree_b200_attention_selection_t* out_selection);

// Records the dispatch(es). |mask| is ignored for MASK_NONE; |workspace| must hold
// selection.workspace_bytes (ignored when 0).
IREE_B200_API iree_b200_status_t iree_b200_ops_record_attention(
    iree_b200_ops_t* ops, iree_b200_command_buffer_t* cb, int32_t element_type, int64_t B,
    int64_t M, int64_t K1, int64_t K2, int64_t N, float scale, int32_t mask_kind, uint32_t flags,
    iree_b200_buffer_ref_t q, iree_b200_buffer_ref_t k, iree_b200_buffer_ref_t v,
    iree_b200_buffer_ref_t mask, iree_b200_buffer_ref_t workspace, iree_b200_buffer_ref_t out);

// Host-buffer form (H2D of Q/K/V/mask, dispatch, D2H of O), synchronous.
IREE_B200_API iree_b200_status_t iree_b200_ops_attention_host(
    iree_b200_ops_t* ops, int32_t element_type, int64_t B, int64_t M, int64_t K1, int64_t K2,
    int64_t N, float scale, int32_t mask_kind, uint32_t flags, const void* q, const void* k,
    const void* v, const void* mask, void* out);

// Split-KV decode attention (K1 == N == 128, no mask; f16 / bf16) for callers that cannot
// allocate. NO COUNTERPART IN THE REFERENCE —
// code under CUDA graph capture, tight serving loops and harnesses that forbid cudaMalloc
// need an entry point that allocates nothing. The partial softmax records the split and merge
// kernels exchange live in a static device array of IREE_B200_NO_ALLOC_SCRATCH_BYTES inside
// the library (every GPU that loads the library holds it until the process exits); the call
// makes no cudaMalloc / cudaFree / allocating library call, takes no ops object and no arena.
//   q [B, M, 128], k, v [B, K2, 128], out [B, M, 128]: DEVICE pointers on |device_ordinal|
//   (primary context), 16-byte aligned; |stream| a CUstream / cudaStream_t of that device
//   (NULL = the default stream). Asynchronous: the two kernels are queued on |stream|.
// Same kernels and split count as the split + merge form of iree_b200_ops_select_attention:
// results are bit-identical to iree_b200_ops_attention_host(..., FLAG_FORCE_DECODE_SPLIT).
// RESOURCE_EXHAUSTED when B * M * splits records of 520 bytes exceed the array (splits =
// min(ceil(K2 / 256), max(1, 8 * SMs / (B * M)))), e.g. B * M > 8065 rows.
// All calls on a device share the one array, which keeps the last call's contents (nothing
// needs clearing: every record read is written first): calls that overlap on the GPU RACE —
// keep them on one stream or order them with events. The first call per device does the
// one-time set-up (cudaGetSymbolAddress): make it before starting a graph capture.
// (iree_b200/csrc/no_alloc/attention_decode_no_alloc.cu)
#define IREE_B200_NO_ALLOC_SCRATCH_BYTES (4u << 20)
IREE_B200_API iree_b200_status_t iree_b200_attention_decode_no_alloc(
    int32_t device_ordinal, int32_t element_type, int64_t B, int64_t M, int64_t K2, float scale,
    const void* q, const void* k, const void* v, void* out, void* stream);

#ifdef __cplusplus
}  // extern "C"
#endif

#endif  // IREE_B200_OPS_H_
