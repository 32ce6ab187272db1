"""iree_b200_attention_decode_no_alloc (include/iree_b200_ops.h): split-KV decode attention whose
scratch is a static device array inside the library, for callers that cannot allocate. No
counterpart in the reference, so the check is
against the existing entry point — iree_b200_ops_attention_host with the split + merge form
forced — and against the oracle, through the C-ABI, on two calls in a row that share the array.

Collected last on purpose. Runs alone in a fresh process or
after the rest of the suite: it brings its own device buffers and touches no shared state but
the static array."""

import numpy as np
import pytest

import oracle
from iree_b200 import hal as H

# The attention tests' bound (tests/test_attention_gpu.py): max|got - want| relative to max|want|.
RTOL_MAX = {"f16": 2e-3, "bf16": 2e-2}


def _qkv(t, B, M, K2, seed):
    g = np.random.Generator(np.random.PCG64(0xA110C + seed))
    mk = lambda *s: oracle.from_f32(t, np.clip(g.standard_normal(s), -4, 4).astype(np.float32))
    return mk(B, M, 128), mk(B, K2, 128), mk(B, K2, 128)


def _no_alloc(device, t, q, k, v, scale):
    """One call of the new entry point on device buffers; returns O as the element type's bits."""
    B, M, _ = q.shape
    K2 = k.shape[1]
    bufs = []
    for a in (q, k, v):
        b = device.allocate(a.nbytes)
        device.h2d(a.ctypes.data, b, a.nbytes)
        bufs.append(b)
    out_b = device.allocate(q.nbytes)
    poison = np.full(q.shape, 0x7e00, dtype=np.uint16)   # a NaN in both f16 and bf16
    device.h2d(poison.ctypes.data, out_b, poison.nbytes)
    device.wait_idle()
    H.attention_decode_no_alloc(device.ordinal, t, B, M, K2, scale, bufs[0].device_pointer,
                                bufs[1].device_pointer, bufs[2].device_pointer,
                                out_b.device_pointer, stream=device.stream)
    device.wait_idle()
    out = np.empty(q.shape, dtype=np.uint16)
    device.d2h(out_b, out.ctypes.data, out.nbytes)
    return out


def _existing(ops, t, q, k, v, scale):
    B, M, _ = q.shape
    out = np.zeros(q.shape, dtype=np.uint16)
    ops.attention_host(t, B, M, 128, k.shape[1], 128, scale, q.ctypes.data, k.ctypes.data,
                       v.ctypes.data, out.ctypes.data,
                       flags=H.ATTENTION_FORCE_DECODE | H.ATTENTION_FORCE_DECODE_SPLIT)
    return out


def _close(t, got, want):
    g, w = oracle.to_f32(t, got).astype(np.float64), oracle.to_f32(t, want).astype(np.float64)
    assert np.isfinite(g).all()
    assert np.abs(g - w).max() <= RTOL_MAX[t] * np.abs(w).max()


@pytest.mark.gpu
@pytest.mark.parametrize("t", ["f16", "bf16"])
def test_no_alloc_decode_matches_the_existing_entry_point_twice_in_a_row(device, ops, t):
    scale = 1.0 / np.sqrt(128.0)
    # Call 1: the Llama decode shape (32 heads x 1 row x 8192 keys: 32 splits per row).
    # Call 2, straight after and differently shaped (ragged K2, M > 1, fewer records): it finds
    # the first call's records in the shared array and must not be affected by them.
    for call, (B, M, K2) in enumerate(((32, 1, 8192), (5, 3, 1000), (32, 1, 8192))):
        q, k, v = _qkv(t, B, M, K2, seed=call)
        got = _no_alloc(device, t, q, k, v, scale)
        existing = _existing(ops, t, q, k, v, scale)
        _close(t, got, existing)
        # Same device code, same split count: not merely close.
        assert np.array_equal(got, existing), f"call {call}: differs from the split + merge export"
        _close(t, got, oracle.attention(t, q, k, v, scale, k2_tile=256))


@pytest.mark.gpu
def test_no_alloc_decode_rejects_what_the_static_array_cannot_hold(device, ops):
    """B * M * splits records of 520 bytes against IREE_B200_NO_ALLOC_SCRATCH_BYTES (4 MiB; on a
    148-SM GPU that is B * M > 8065 rows): a clear RESOURCE_EXHAUSTED, nothing launched, and the
    next call still works."""
    dummy = device.allocate(4096).device_pointer
    for B, M, K2 in ((9000, 1, 256), (8066, 1, 8192), (130, 63, 1024)):
        assert B * M * 520 > H.NO_ALLOC_SCRATCH_BYTES
        with pytest.raises(H.Status) as e:
            H.attention_decode_no_alloc(device.ordinal, "f16", B, M, K2, 0.1, dummy, dummy, dummy,
                                        dummy, stream=device.stream)
        assert e.value.name == "RESOURCE_EXHAUSTED"
        assert "IREE_B200_NO_ALLOC_SCRATCH_BYTES" in e.value.message
    for bad in (dict(t="f32"), dict(K2=0), dict(q=dummy + 8)):
        with pytest.raises(H.Status) as e:
            H.attention_decode_no_alloc(device.ordinal, bad.get("t", "f16"), 4, 1, bad.get("K2", 512),
                                        0.1, bad.get("q", dummy), dummy, dummy, dummy)
        assert e.value.name == "INVALID_ARGUMENT"
    H.attention_decode_no_alloc(device.ordinal, "f16", 0, 1, 512, 0.1, 0, 0, 0, 0)   # empty: no-op
    scale = 1.0 / np.sqrt(128.0)
    q, k, v = _qkv("f16", 8, 2, 700, seed=9)
    assert np.array_equal(_no_alloc(device, "f16", q, k, v, scale), _existing(ops, "f16", q, k, v, scale))
