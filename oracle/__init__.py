"""ctypes wrapper over oracle/liboracle.so — the CPU restatement of the reference's
arithmetic for matmul / batch_matmul / attention.

TEST INFRASTRUCTURE ONLY. Importable from tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs; never from iree_b200/.
See oracle/oracle.c for the per-function reference citations and parity status.
"""

from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")

I8, F16, BF16, F32, I32, F8E4M3FN, F8E5M2, U8, F8E8M0FNU = 0, 1, 2, 3, 4, 5, 6, 7, 8
_TYPE_BY_NAME = {"i8": I8, "f16": F16, "bf16": BF16, "f32": F32, "i32": I32,
                 "f8e4m3fn": F8E4M3FN, "f8e5m2": F8E5M2, "u8": U8, "f8e8m0fnu": F8E8M0FNU}
_NP_STORAGE = {I8: np.int8, F16: np.uint16, BF16: np.uint16, F32: np.float32, I32: np.int32,
               F8E4M3FN: np.uint8, F8E5M2: np.uint8, U8: np.uint8, F8E8M0FNU: np.uint8}


def build(force: bool = False) -> str:
    """Compiles liboracle.so (gcc, OpenMP). Building the checker is not using it."""
    src = os.path.join(_HERE, "oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.run(["make", "-C", _HERE, "liboracle.so"], check=True,
                       stdout=subprocess.DEVNULL)
    return _LIB_PATH


_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_LIB_PATH)
        _lib.oracle_bf16_to_f32.restype = ctypes.c_float
        _lib.oracle_f16_to_f32.restype = ctypes.c_float
        _lib.oracle_f32_to_bf16.restype = ctypes.c_uint16
        _lib.oracle_f32_to_f16.restype = ctypes.c_uint16
        _lib.oracle_f32_to_bf16.argtypes = [ctypes.c_float]
        _lib.oracle_f32_to_f16.argtypes = [ctypes.c_float]
        for name in ("oracle_f8e4m3fn_to_f32", "oracle_f8e5m2_to_f32", "oracle_f4e2m1fn_to_f32",
                     "oracle_f8e8m0fnu_to_f32"):
            getattr(_lib, name).restype = ctypes.c_float
            getattr(_lib, name).argtypes = [ctypes.c_uint8]
        for name in ("oracle_f32_to_f8e4m3fn", "oracle_f32_to_f8e5m2", "oracle_f32_to_f8e8m0fnu"):
            getattr(_lib, name).restype = ctypes.c_uint8
            getattr(_lib, name).argtypes = [ctypes.c_float]
        _lib.oracle_matmul_f32acc_element.restype = ctypes.c_float
        _lib.oracle_matmul_i8_i32_element.restype = ctypes.c_int32
        _lib.oracle_calculate_check_every.restype = ctypes.c_int
    return _lib


def type_id(t) -> int:
    return _TYPE_BY_NAME[t] if isinstance(t, str) else int(t)


def _ptr(a):
    if a is None:
        return ctypes.c_void_p(0)
    assert a.flags["C_CONTIGUOUS"]
    return ctypes.c_void_p(a.ctypes.data)


# ----------------------------------------------------------------------------
# bf16 / f16 bit helpers (numpy has no bf16; narrow floats travel as uint16 bits).
# ----------------------------------------------------------------------------

def f32_to_bf16_bits(x: np.ndarray) -> np.ndarray:
    u = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32).astype(np.uint64)
    nan = (u & 0x7F800000) == 0x7F800000
    r = ((u + 0x7FFF + ((u >> 16) & 1)) >> 16).astype(np.uint16)
    r = np.where(nan, (u >> 16).astype(np.uint16), r)
    return r


def bf16_bits_to_f32(b: np.ndarray) -> np.ndarray:
    return (b.astype(np.uint32) << 16).view(np.float32)


def f32_to_f16_bits(x: np.ndarray) -> np.ndarray:
    return np.ascontiguousarray(x, dtype=np.float32).astype(np.float16).view(np.uint16)


def f16_bits_to_f32(b: np.ndarray) -> np.ndarray:
    return b.view(np.float16).astype(np.float32)


_F8_NAMES = {F8E4M3FN: "f8e4m3fn", F8E5M2: "f8e5m2"}
_f8_tables = {}


def _f8_decode_table(t) -> np.ndarray:
    """All 256 codes of an 8-bit float decoded by the C oracle (math.h:300-362 restated)."""
    if t not in _f8_tables:
        fn = getattr(lib(), "oracle_%s_to_f32" % _F8_NAMES[t])
        _f8_tables[t] = np.array([fn(i) for i in range(256)], dtype=np.float32)
    return _f8_tables[t]


def to_f32(t, a: np.ndarray) -> np.ndarray:
    t = type_id(t)
    if t == BF16:
        return bf16_bits_to_f32(a)
    if t == F16:
        return f16_bits_to_f32(a)
    if t in _F8_NAMES:
        return _f8_decode_table(t)[a]
    return a.astype(np.float32)


def from_f32(t, a: np.ndarray) -> np.ndarray:
    t = type_id(t)
    if t == BF16:
        return f32_to_bf16_bits(a)
    if t == F16:
        return f32_to_f16_bits(a)
    if t in _F8_NAMES:
        # Round to nearest even through the C encoder, once per distinct value.
        fn = getattr(lib(), "oracle_f32_to_%s" % _F8_NAMES[t])
        flat = np.ascontiguousarray(a, dtype=np.float32).reshape(-1)
        uniq, inverse = np.unique(flat.view(np.uint32), return_inverse=True)
        codes = np.array([fn(float(u)) for u in uniq.view(np.float32)], dtype=np.uint8)
        return codes[inverse].reshape(np.shape(a))
    return np.ascontiguousarray(a, dtype=np.float32)


# ----------------------------------------------------------------------------
# Reference input generator.
# ----------------------------------------------------------------------------

def generate_random_matrix(t, shape, seed: int) -> np.ndarray:
    """iree-e2e-matmul-test.cc:694-736 GenerateRandomMatrix (see oracle.c)."""
    t = type_id(t)
    out = np.empty(shape, dtype=_NP_STORAGE[t])
    lib().oracle_generate_random_matrix(ctypes.c_int(t), ctypes.c_int64(out.size),
                                        ctypes.c_int32(seed), _ptr(out))
    return out


def calculate_check_every(tot_elements: int, no_div_of: int, max_elements: int = 10000) -> int:
    return lib().oracle_calculate_check_every(ctypes.c_int64(tot_elements),
                                              ctypes.c_int64(no_div_of),
                                              ctypes.c_int32(max_elements))


# ----------------------------------------------------------------------------
# matmul / batch_matmul
# ----------------------------------------------------------------------------

def matmul(t, lhs: np.ndarray, rhs: np.ndarray, acc=None, transpose_rhs: bool = False) -> np.ndarray:
    """C = acc + A.B with the reference's per-element k-ascending accumulation.

    lhs [.., M, K]; rhs [.., K, N] (or [.., N, K] when transpose_rhs). A leading
    batch dim makes it batch_matmul. t in {"i8","f16","bf16","f8e4m3fn","f8e5m2"};
    result i32 / f32.
    """
    t = type_id(t)
    batched = lhs.ndim == 3
    B = lhs.shape[0] if batched else 1
    M, K = lhs.shape[-2:]
    N = rhs.shape[-2] if transpose_rhs else rhs.shape[-1]
    lhs = np.ascontiguousarray(lhs)
    rhs = np.ascontiguousarray(rhs)
    oshape = (B, M, N) if batched else (M, N)
    if t == I8:
        out = np.empty(oshape, dtype=np.int32)
        acc_c = None if acc is None else np.ascontiguousarray(acc, dtype=np.int32)
        lib().oracle_matmul_i8_i32(ctypes.c_int64(B), ctypes.c_int64(M), ctypes.c_int64(N),
                                   ctypes.c_int64(K), ctypes.c_int(int(transpose_rhs)),
                                   _ptr(lhs), _ptr(rhs), _ptr(acc_c), _ptr(out))
    else:
        out = np.empty(oshape, dtype=np.float32)
        acc_c = None if acc is None else np.ascontiguousarray(acc, dtype=np.float32)
        lib().oracle_matmul_f32acc(ctypes.c_int(t), ctypes.c_int64(B), ctypes.c_int64(M),
                                   ctypes.c_int64(N), ctypes.c_int64(K),
                                   ctypes.c_int(int(transpose_rhs)), _ptr(lhs), _ptr(rhs),
                                   _ptr(acc_c), _ptr(out))
    return out


def matmul_bias(t, lhs: np.ndarray, rhs: np.ndarray, bias: np.ndarray, acc=None,
                transpose_rhs: bool = False, relu: bool = False, truncate: bool = False) -> np.ndarray:
    """The fused matmul + elementwise-consumer dispatch, op by op as the unfused IR evaluates it:
    linalg.matmul (above), then linalg.generic { arith.addf / addi  %mm, %bias[n] }
    [ arith.maximumf / maxsi %x, 0 ] [ arith.truncf ]. Each op rounds once in f32 (i32 wraps);
    maximumf propagates NaN and returns +0 for (-0, +0)."""
    x = matmul(t, lhs, rhs, acc=acc, transpose_rhs=transpose_rhs)
    bias = np.asarray(bias, dtype=x.dtype)
    assert bias.shape == (x.shape[-1],)
    with np.errstate(over="ignore", invalid="ignore"):
        x = (x + bias).astype(x.dtype)
        if relu:
            keep = x > 0
            if x.dtype == np.float32:
                keep |= np.isnan(x)
            x = np.where(keep, x, x.dtype.type(0))
    return from_f32(t, x) if truncate else x


MX_BLOCK = 32


def mx_decode(bytes_: np.ndarray, scales: np.ndarray, block: int = MX_BLOCK) -> np.ndarray:
    """arith.scaling_extf of packed f4E2M1FN rows [R, K/2] (low nibble first) by their
    f8E8M0FNU block scales [R, K/block] -> f32 [R, K]."""
    f4 = np.array([lib().oracle_f4e2m1fn_to_f32(i) for i in range(16)], dtype=np.float32)
    e8 = np.array([lib().oracle_f8e8m0fnu_to_f32(i) for i in range(256)], dtype=np.float32)
    b = np.ascontiguousarray(bytes_, dtype=np.uint8)
    codes = np.stack([b & 0xF, b >> 4], axis=-1).reshape(b.shape[0], -1)
    with np.errstate(invalid="ignore", over="ignore"):
        return f4[codes] * np.repeat(e8[scales], block, axis=1)


def matmul_mxfp4(lhs: np.ndarray, lhs_scales: np.ndarray, rhs: np.ndarray, rhs_scales: np.ndarray,
                 acc=None, block: int = MX_BLOCK) -> np.ndarray:
    """MXFP4 matmul as tests/e2e/matmul/generate_code_mx.py states it (see oracle.c):
    lhs [M, K/2] / rhs [N, K/2] packed f4E2M1FN bytes, scales [M|N, K/block] f8E8M0FNU,
    result f32 [M, N]."""
    lhs, rhs = (np.ascontiguousarray(x, dtype=np.uint8) for x in (lhs, rhs))
    lhs_scales, rhs_scales = (np.ascontiguousarray(x, dtype=np.uint8)
                              for x in (lhs_scales, rhs_scales))
    M, K = lhs.shape[0], lhs.shape[1] * 2
    N = rhs.shape[0]
    assert rhs.shape[1] * 2 == K and K % block == 0
    assert lhs_scales.shape == (M, K // block) and rhs_scales.shape == (N, K // block)
    out = np.empty((M, N), dtype=np.float32)
    acc_c = None if acc is None else np.ascontiguousarray(acc, dtype=np.float32)
    lib().oracle_matmul_mxfp4_f32(ctypes.c_int64(M), ctypes.c_int64(N), ctypes.c_int64(K),
                                  ctypes.c_int64(block), _ptr(lhs), _ptr(lhs_scales), _ptr(rhs),
                                  _ptr(rhs_scales), _ptr(acc_c), _ptr(out))
    return out


def pack(a: np.ndarray, tile0: int, tile1: int) -> np.ndarray:
    """[R, C] -> [ceil(R/tile0), ceil(C/tile1), tile0, tile1], zero padded: linalg.pack with
    inner_dims_pos = [0, 1] (what iree_uk_pack produces for the LHS; the RHS of a mmt4d is
    pack(rhs.T, N0, K0))."""
    R, C = a.shape
    R1, C1 = -(-R // tile0), -(-C // tile1)
    padded = np.zeros((R1 * tile0, C1 * tile1), dtype=a.dtype)
    padded[:R, :C] = a
    return np.ascontiguousarray(padded.reshape(R1, tile0, C1, tile1).transpose(0, 2, 1, 3))


def unpack(a4: np.ndarray, rows: int, cols: int) -> np.ndarray:
    """Inverse of pack(): [R1, C1, tile0, tile1] -> [rows, cols]."""
    R1, C1, t0, t1 = a4.shape
    return np.ascontiguousarray(a4.transpose(0, 2, 1, 3).reshape(R1 * t0, C1 * t1)[:rows, :cols])


def mmt4d(t, lhs: np.ndarray, rhs: np.ndarray, acc=None) -> np.ndarray:
    """linalg.mmt4d (see oracle.c): lhs [M1,K1,M0,K0], rhs [N1,K1,N0,K0] -> [M1,N1,M0,N0]."""
    tid = type_id(t)
    lhs, rhs = np.ascontiguousarray(lhs), np.ascontiguousarray(rhs)
    M1, K1, M0, K0 = lhs.shape
    N1, K1r, N0, K0r = rhs.shape
    assert (K1, K0) == (K1r, K0r)
    out_t = np.int32 if t == "i8" else np.float32
    out = np.empty((M1, N1, M0, N0), dtype=out_t)
    acc_c = None if acc is None else np.ascontiguousarray(acc, dtype=out_t)
    assert acc_c is None or acc_c.shape == out.shape
    rc = lib().oracle_mmt4d(ctypes.c_int(tid), *(ctypes.c_int64(v) for v in (M1, N1, K1, M0, N0, K0)),
                            _ptr(lhs), _ptr(rhs), _ptr(acc_c), _ptr(out))
    if rc != 0:
        raise ValueError(f"mmt4d: unsupported element type {t}")
    return out


def ref_mmt4d(t, lhs: np.ndarray, rhs: np.ndarray, acc=None, force_generic: bool = False) -> np.ndarray:
    """The same contraction through the reference's own iree_uk_mmt4d (oracle/_ref)."""
    tid = type_id(t)
    lhs, rhs = np.ascontiguousarray(lhs), np.ascontiguousarray(rhs)
    M1, K1, M0, K0 = lhs.shape
    N1, _, N0, _ = rhs.shape
    out_t = np.int32 if t == "i8" else np.float32
    out = (np.zeros((M1, N1, M0, N0), dtype=out_t) if acc is None
           else np.ascontiguousarray(acc, dtype=out_t).copy())
    fn = ref_lib().oracle_ref_mmt4d
    fn.restype = ctypes.c_int
    rc = fn(ctypes.c_int(tid), ctypes.c_int64(M1), ctypes.c_int64(N1), ctypes.c_int64(K1),
            ctypes.c_int(M0), ctypes.c_int(N0), ctypes.c_int(K0), _ptr(lhs), _ptr(rhs), _ptr(out),
            ctypes.c_int(int(acc is not None)), ctypes.c_int(int(force_generic)))
    if rc != 0:
        raise RuntimeError("oracle_ref_mmt4d failed")
    return out


def matmul_element(t, lhs: np.ndarray, rhs: np.ndarray, m: int, n: int, transpose_rhs: bool = False):
    t = type_id(t)
    M, K = lhs.shape
    N = rhs.shape[0] if transpose_rhs else rhs.shape[1]
    args = [ctypes.c_int64(M), ctypes.c_int64(N), ctypes.c_int64(K),
            ctypes.c_int(int(transpose_rhs)), _ptr(lhs), _ptr(rhs),
            ctypes.c_int64(m), ctypes.c_int64(n)]
    if t == I8:
        return lib().oracle_matmul_i8_i32_element(*args)
    return lib().oracle_matmul_f32acc_element(ctypes.c_int(t), *args)


def sampled_indices(M: int, N: int, max_elements: int = 10000):
    """Flat indices the reference's checker visits (test_utils.c:41-68)."""
    every = calculate_check_every(M * N, N, max_elements)
    return np.arange(0, M * N, every, dtype=np.int64), every


# ----------------------------------------------------------------------------
# attention
# ----------------------------------------------------------------------------

def attention(t, q: np.ndarray, k: np.ndarray, v: np.ndarray, scale: float, *, mask=None,
              scale_type=None, region: str = "identity", region_param: float = 0.0,
              k2_tile: int = 0, oneshot: bool = False) -> np.ndarray:
    """iree_linalg_ext.attention as the llvm-cpu path evaluates it (see oracle.c).

    q [B,M,K1], k [B,K2,K1], v [B,K2,N] in storage dtype of t (uint16 bits for
    f16/bf16, float32 for f32). mask: None | bool [B,M,K2] | float32 [B,M,K2].
    """
    t = type_id(t)
    st = t if scale_type is None else type_id(scale_type)
    B, M, K1 = q.shape
    K2 = k.shape[1]
    N = v.shape[2]
    q, k, v = (np.ascontiguousarray(x) for x in (q, k, v))
    out = np.empty((B, M, N), dtype=_NP_STORAGE[t])
    mask_kind, mask_c = 0, None
    if mask is not None:
        if mask.dtype == np.bool_:
            mask_kind, mask_c = 1, np.ascontiguousarray(mask, dtype=np.uint8)
        else:
            mask_kind, mask_c = 2, np.ascontiguousarray(mask, dtype=np.float32)
        assert mask_c.shape == (B, M, K2)
    region_kind = {"identity": 0, "softcap": 1}[region]
    common = [ctypes.c_int(t), ctypes.c_int(st), ctypes.c_int64(B), ctypes.c_int64(M),
              ctypes.c_int64(K1), ctypes.c_int64(K2), ctypes.c_int64(N), _ptr(q), _ptr(k),
              _ptr(v), ctypes.c_float(scale), ctypes.c_int(mask_kind), _ptr(mask_c),
              ctypes.c_int(region_kind), ctypes.c_float(region_param)]
    if oneshot:
        lib().oracle_attention_oneshot(*common, _ptr(out))
    else:
        lib().oracle_attention_online(*common, ctypes.c_int64(k2_tile), _ptr(out))
    return out


def all_reduce_sum_f32(parts: np.ndarray) -> np.ndarray:
    parts = np.ascontiguousarray(parts, dtype=np.float32)
    out = np.empty(parts.shape[1:], dtype=np.float32)
    lib().oracle_all_reduce_sum_f32(ctypes.c_int(parts.shape[0]),
                                    ctypes.c_int64(out.size), _ptr(parts), _ptr(out))
    return out


# ----------------------------------------------------------------------------
# oracle/_ref: the reference's own CPU micro-kernels (built by oracle/build_ref.sh)
# ----------------------------------------------------------------------------

_REF_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "libiree_uk_ref.so")
_ref_lib = None


def ref_available() -> bool:
    return os.path.exists(_REF_PATH)


def ref_lib() -> ctypes.CDLL:
    """oracle/_ref/libiree_uk_ref.so: iree_uk_{query_tile_sizes,pack,mmt4d,unpack} compiled
    from /root/reference plus oracle/ref_glue.c. Built where /root/reference exists
    (`make -C oracle ref`); the .so travels to the GPU box."""
    global _ref_lib
    if _ref_lib is None:
        l = ctypes.CDLL(_REF_PATH)
        l.oracle_ref_matmul.restype = ctypes.c_int
        l.oracle_ref_matmul.argtypes = [ctypes.c_int, ctypes.c_int64, ctypes.c_int64,
                                        ctypes.c_int64, ctypes.c_void_p, ctypes.c_void_p,
                                        ctypes.c_void_p, ctypes.c_int,
                                        ctypes.POINTER(ctypes.c_int)]
        l.oracle_ref_cpu_features.restype = ctypes.c_uint64
        _ref_lib = l
    return _ref_lib


def ref_matmul(t, lhs: np.ndarray, rhs: np.ndarray, force_generic: bool = False,
               return_tiles: bool = False):
    """Row-major [M,K] x [K,N] through the reference ukernels (pack -> mmt4d -> unpack)."""
    tid = type_id(t)
    M, K = lhs.shape
    N = rhs.shape[1]
    lhs, rhs = np.ascontiguousarray(lhs), np.ascontiguousarray(rhs)
    out = np.empty((M, N), dtype=np.int32 if t == "i8" else np.float32)
    tiles = (ctypes.c_int * 3)()
    rc = ref_lib().oracle_ref_matmul(tid, M, N, K, _ptr(lhs), _ptr(rhs), _ptr(out),
                                     int(force_generic), tiles)
    if rc != 0:
        raise RuntimeError(f"oracle_ref_matmul failed for {t} {M}x{N}x{K}")
    return (out, tuple(tiles)) if return_tiles else out


_ENCODE_KINDS = {"f8e4m3fn": 0, "f8e5m2": 1, "f16": 2, "bf16": 3, "f8e8m0fnu": 4}


def encode_bulk(t: str, values: np.ndarray, reference: bool = False) -> np.ndarray:
    """f32 -> narrow float bits for every element: oracle.c's restatement, or — with
    |reference| — the reference's own iree_math_f32_to_* (math.h, via oracle/_ref)."""
    v = np.ascontiguousarray(values, dtype=np.float32).reshape(-1)
    out = np.empty(v.size, dtype=np.uint16)
    fn = ref_lib().oracle_ref_encode if reference else lib().oracle_encode
    fn(ctypes.c_int(_ENCODE_KINDS[t]), ctypes.c_int64(v.size), _ptr(v), _ptr(out))
    return out


def decode_reference(t: str, bits: np.ndarray) -> np.ndarray:
    """narrow float bits -> f32 through the reference's iree_math_*_to_f32 (oracle/_ref)."""
    fn = getattr(ref_lib(), "oracle_ref_%s_to_f32" % t)
    fn.restype = ctypes.c_float
    fn.argtypes = [ctypes.c_uint8 if t.startswith(("f8", "f4")) else ctypes.c_uint16]
    return np.array([fn(int(b)) for b in np.asarray(bits).reshape(-1)], dtype=np.float32)
