"""Conformance of the thin launch layer (iree_b200/csrc/hal) on a B200, in the pattern of
the reference's HAL CTS (runtime/src/iree/hal/cts/{buffer,command_buffer,core,queue}/*.cc):
allocator + transfers, fill / update / copy, barriers, dispatch with constants and bindings,
dispatch reuse, executable cache, semaphores and queue ordering, recording state machine —
each against BOTH command-buffer implementations (STREAM replay and CUgraph), and the
tcgen05 / attention dispatches of the ops layer replayed as graphs.
"""

import ctypes
import struct

import numpy as np
import pytest

import oracle
from iree_b200 import hal as H

pytestmark = pytest.mark.gpu

MODES = {"stream": H.MODE_STREAM, "graph": H.MODE_GRAPH}


@pytest.fixture(scope="module", params=list(MODES), ids=list(MODES))
def dev(request):
    d = H.Device(0, mode=MODES[request.param])
    d._mode_name = request.param
    return d


@pytest.fixture(scope="module")
def dev_ops(dev):
    return H.Ops(dev)


def _upload(dev, array):
    array = np.ascontiguousarray(array)
    b = dev.allocate(array.nbytes)
    dev.h2d(array.ctypes.data, b, array.nbytes)
    return b


def _download(dev, buffer, dtype, shape, offset=0):
    out = np.empty(shape, dtype=dtype)
    dev.d2h(buffer, out.ctypes.data, out.nbytes, offset=offset)
    return out


def _run(dev, record):
    cb = H.CommandBuffer(dev).begin()
    record(cb)
    cb.end()
    dev.execute(cb)
    dev.wait_idle()
    return cb


# ---- buffer/allocator_test.cc, mapping_test.cc --------------------------------------

def test_allocator_and_transfers(dev):
    b = dev.allocate(4096)
    assert b.byte_length == 4096 and b.device_pointer and b.device_pointer % 256 == 0
    src = np.arange(1024, dtype=np.uint32)
    dev.h2d(src.ctypes.data, b, 4096)
    assert np.array_equal(_download(dev, b, np.uint32, 1024), src)
    # Sub-range transfers leave the rest alone.
    patch = np.full(16, 0xABCD1234, dtype=np.uint32)
    dev.h2d(patch.ctypes.data, b, 64, offset=128)
    want = src.copy()
    want[32:48] = patch
    assert np.array_equal(_download(dev, b, np.uint32, 1024), want)
    assert np.array_equal(_download(dev, b, np.uint32, 16, offset=128), patch)
    with pytest.raises(H.Status) as e:      # past the end
        dev.h2d(patch.ctypes.data, b, 64, offset=4096 - 32)
    assert e.value.name == "OUT_OF_RANGE"
    assert dev.allocate(0).byte_length == 0  # zero-length buffers are legal
    with pytest.raises(H.Status):
        dev.allocate(64, memory_type=77)


def test_host_local_buffers_are_mapped(dev):
    b = dev.allocate(256, memory_type=H.MEMORY_HOST_LOCAL)
    host = H.lib().iree_b200_buffer_host_pointer(b._h)
    assert host and b.device_pointer
    view = (ctypes.c_uint8 * 256).from_address(host)
    for i in range(256):
        view[i] = i
    # The device sees the mapping: copy it into device memory with a command buffer.
    d = dev.allocate(256)
    _run(dev, lambda cb: cb.copy(b.ref(), d.ref()))
    assert np.array_equal(_download(dev, d, np.uint8, 256), np.arange(256, dtype=np.uint8))
    # ... and device writes land in the host view.
    _run(dev, lambda cb: cb.fill(b.ref(16, 32), b"\xee"))
    assert bytes(view[16:48]) == b"\xee" * 32 and view[15] == 15 and view[48] == 48
    assert H.lib().iree_b200_buffer_host_pointer(d._h) in (None, 0)


def test_imported_buffers_alias_the_allocation(dev):
    owner = dev.allocate(1024)
    alias = dev.import_buffer(owner.device_pointer + 256, 512)
    _run(dev, lambda cb: cb.fill(owner.ref(), b"\x00"))
    _run(dev, lambda cb: cb.fill(alias.ref(), b"\x5a"))
    got = _download(dev, owner, np.uint8, 1024)
    assert (got[:256] == 0).all() and (got[256:768] == 0x5A).all() and (got[768:] == 0).all()
    with pytest.raises(H.Status):
        dev.import_buffer(0, 16)


# ---- command_buffer/{fill,update,copy}_buffer_test.cc ------------------------------------

@pytest.mark.parametrize("pattern", [b"\x07", b"\x34\x12", b"\xef\xbe\xad\xde"], ids=["u8", "u16", "u32"])
def test_fill_buffer(dev, pattern):
    n = len(pattern)
    b = _upload(dev, np.zeros(256, dtype=np.uint8))
    _run(dev, lambda cb: cb.fill(b.ref(16 * n, 32 * n), pattern))
    got = _download(dev, b, np.uint8, 256).tobytes()
    assert got == b"\x00" * (16 * n) + pattern * 32 + b"\x00" * (256 - 48 * n)
    # Whole-buffer form and an empty range.
    _run(dev, lambda cb: (cb.fill(b.ref(), pattern), cb.fill(b.ref(8, 0), pattern)))
    assert _download(dev, b, np.uint8, 256).tobytes() == pattern * (256 // n)


def test_fill_buffer_guards(dev):
    b = dev.allocate(64)
    cb = H.CommandBuffer(dev).begin()
    with pytest.raises(H.Status) as e:
        cb.fill(b.ref(), b"\x01\x02\x03")            # pattern length 3
    assert e.value.name == "INVALID_ARGUMENT"
    with pytest.raises(H.Status):
        cb.fill(b.ref(2, 8), b"\x01\x02\x03\x04")    # offset not a multiple of the pattern
    with pytest.raises(H.Status):
        cb.fill(b.ref(0, 6), b"\x01\x02\x03\x04")    # length not a multiple of the pattern
    with pytest.raises(H.Status) as e:
        cb.fill(b.ref(32, 64), b"\x01")              # past the end
    assert e.value.name == "OUT_OF_RANGE"


def test_update_buffer_captures_its_source(dev):
    b = _upload(dev, np.zeros(64, dtype=np.uint8))
    staging = (ctypes.c_uint8 * 16)(*range(1, 17))
    cb = H.CommandBuffer(dev).begin()
    cb.update(staging, b.ref(8, 16))
    for i in range(16):
        staging[i] = 0xFF                             # must not be seen: the bytes were captured
    cb.end()
    dev.execute(cb)
    dev.wait_idle()
    got = _download(dev, b, np.uint8, 64)
    assert np.array_equal(got[8:24], np.arange(1, 17, dtype=np.uint8))
    assert (got[:8] == 0).all() and (got[24:] == 0).all()


def test_copy_buffer(dev):
    src = _upload(dev, np.arange(256, dtype=np.uint8))
    dst = _upload(dev, np.zeros(256, dtype=np.uint8))
    _run(dev, lambda cb: cb.copy(src.ref(100, 50), dst.ref(7, 50)))
    got = _download(dev, dst, np.uint8, 256)
    assert np.array_equal(got[7:57], np.arange(100, 150, dtype=np.uint8))
    assert (got[:7] == 0).all() and (got[57:] == 0).all()
    cb = H.CommandBuffer(dev).begin()
    with pytest.raises(H.Status) as e:
        cb.copy(src.ref(0, 16), dst.ref(0, 32))      # lengths differ
    assert e.value.name == "INVALID_ARGUMENT"
    with pytest.raises(H.Status) as e:
        cb.copy(src.ref(250, 16), dst.ref(0, 16))    # source range past the end
    assert e.value.name == "OUT_OF_RANGE"


def test_execution_barriers_order_dependent_commands(dev):
    """fill A | copy A->B | refill A | copy A->C: every stage depends on the one before
    (graph mode: the barriers are the only edges, graph_command_buffer.c:413-439)."""
    a, b, c = (_upload(dev, np.zeros(1 << 20, dtype=np.uint8)) for _ in range(3))

    def record(cb):
        cb.fill(a.ref(), b"\x11")
        cb.barrier()
        cb.copy(a.ref(), b.ref())
        cb.barrier()
        cb.fill(a.ref(), b"\x22")
        cb.barrier()
        cb.copy(a.ref(), c.ref())

    cb = _run(dev, record)
    assert (_download(dev, b, np.uint8, 1 << 20) == 0x11).all()
    assert (_download(dev, c, np.uint8, 1 << 20) == 0x22).all()
    # Replay (dispatch_reuse_test.cc): same result, no re-recording.
    _run(dev, lambda cb2: (cb2.fill(b.ref(), b"\x00"), cb2.fill(c.ref(), b"\x00")))
    dev.execute(cb)
    dev.wait_idle()
    assert (_download(dev, b, np.uint8, 1 << 20) == 0x11).all()
    assert (_download(dev, c, np.uint8, 1 << 20) == 0x22).all()
    # A barrier-only and an empty command buffer are fine too.
    _run(dev, lambda cb3: cb3.barrier())
    _run(dev, lambda cb4: None)


# ---- command_buffer/dispatch*_test.cc ---------------------------------------------------

def _generic_i8(dev_ops):
    exe = dev_ops.executable
    ordinal = exe.lookup("iree_b200_matmul_generic_i8_i32")
    info = exe.export_info(ordinal)
    assert info["constant_count"] == 5 and info["binding_count"] == 3
    return exe, ordinal


def test_dispatch_constants_and_binding_offsets(dev, dev_ops):
    """Raw dispatch of one export: constants in order, binding pointer = base + offset
    (dispatch_constants_bindings_test.cc; kernels.cu:28-34)."""
    exe, ordinal = _generic_i8(dev_ops)
    M, N, K = 70, 50, 33
    lhs = oracle.generate_random_matrix("i8", (M, K), 1)
    rhs = oracle.generate_random_matrix("i8", (K, N), 2)
    # All three operands live in ONE buffer at unaligned (element-aligned) offsets.
    off_l, off_r = 3, 3 + M * K + 5
    off_o = (off_r + K * N + 3) // 4 * 4 + 4
    blob = np.zeros(off_o + M * N * 4 + 8, dtype=np.uint8)
    blob[off_l:off_l + M * K] = lhs.view(np.uint8).ravel()
    blob[off_r:off_r + K * N] = rhs.view(np.uint8).ravel()
    b = _upload(dev, blob)
    grid = [-(-N // 64), -(-M // 64), 1]
    _run(dev, lambda cb: cb.dispatch(exe, ordinal, grid, [M, N, K, 1, 0],
                                     [b.ref(off_l, M * K), b.ref(off_r, K * N),
                                      b.ref(off_o, M * N * 4)]))
    got = _download(dev, b, np.uint8, blob.size)
    assert np.array_equal(got[off_o:off_o + M * N * 4].view(np.int32).reshape(M, N),
                          oracle.matmul("i8", lhs, rhs))
    assert np.array_equal(got[:off_o], blob[:off_o]) and (got[off_o + M * N * 4:] == 0).all()


def test_dispatch_argument_checks(dev, dev_ops):
    exe, ordinal = _generic_i8(dev_ops)
    b = dev.allocate(1024)
    refs = [b.ref(0, 64), b.ref(64, 64), b.ref(128, 256)]
    cb = H.CommandBuffer(dev).begin()
    with pytest.raises(H.Status) as e:
        cb.dispatch(exe, ordinal, [1, 1, 1], [8, 8, 8, 1], refs)          # 4 of 5 constants
    assert e.value.name == "INVALID_ARGUMENT" and "constants" in e.value.message
    with pytest.raises(H.Status) as e:
        cb.dispatch(exe, ordinal, [1, 1, 1], [8, 8, 8, 1, 0], refs[:2])   # 2 of 3 bindings
    assert e.value.name == "INVALID_ARGUMENT" and "bindings" in e.value.message
    with pytest.raises(H.Status) as e:
        cb.dispatch(exe, exe.export_count, [1, 1, 1], [8, 8, 8, 1, 0], refs)
    assert e.value.name == "OUT_OF_RANGE"
    with pytest.raises(H.Status) as e:
        cb.dispatch(exe, ordinal, [1, 1, 1], [8, 8, 8, 1, 0], [refs[0], refs[1], b.ref(1000, 64)])
    assert e.value.name == "OUT_OF_RANGE"
    # A zero-sized grid records nothing and is not an error.
    cb.dispatch(exe, ordinal, [0, 1, 1], [8, 8, 8, 1, 0], refs)
    cb.end()
    dev.execute(cb)
    dev.wait_idle()


def test_dispatch_reuse_accumulates(dev, dev_ops):
    """The same executable command buffer submitted three times (dispatch_reuse_test.cc):
    with the accumulate flag each run adds A.B once more."""
    exe, ordinal = _generic_i8(dev_ops)
    M = N = K = 64
    lhs = oracle.generate_random_matrix("i8", (M, K), 1)
    rhs = oracle.generate_random_matrix("i8", (K, N), 2)
    bl, br = _upload(dev, lhs), _upload(dev, rhs)
    bo = _upload(dev, np.zeros((M, N), dtype=np.int32))
    cb = H.CommandBuffer(dev).begin()
    cb.dispatch(exe, ordinal, [1, 1, 1], [M, N, K, 1, 1], [bl.ref(), br.ref(), bo.ref()])
    cb.end()
    for _ in range(3):
        dev.execute(cb)
    dev.wait_idle()
    assert np.array_equal(_download(dev, bo, np.int32, (M, N)), 3 * oracle.matmul("i8", lhs, rhs))


@pytest.mark.parametrize("t", ["bf16", "i8", "f8e4m3fn"])
@pytest.mark.parametrize("family", ["2cta", "1cta"])
def test_tma_dispatches_in_both_command_buffer_kinds(dev, dev_ops, t, family):
    """The TMA-fed exports (trailing CUtensorMap parameters, compiled-in cluster dims,
    >48 KiB dynamic shared memory) as stream launches and as CUgraph kernel nodes."""
    M, N, K = 384, 512, 256
    lhs = oracle.generate_random_matrix(t, (M, K), 1)
    rhs = oracle.generate_random_matrix(t, (K, N), 2)
    bl, br = _upload(dev, lhs), _upload(dev, rhs)
    out_dtype = np.int32 if t == "i8" else np.float32
    bo = _upload(dev, np.full((M, N), 77, dtype=out_dtype))
    flag = {"2cta": H.MATMUL_FORCE_2CTA, "1cta": H.MATMUL_FORCE_1CTA}[family]
    cb = H.CommandBuffer(dev).begin()
    dev_ops.record_matmul(cb, t, 1, M, N, K, bl.ref(), br.ref(), bo.ref(), flags=flag)
    cb.end()
    want = oracle.matmul(t, lhs, rhs)
    for _ in range(2):     # replayed: the encoded tensor maps live in the command buffer
        dev.execute(cb)
        dev.wait_idle()
        assert np.array_equal(_download(dev, bo, out_dtype, (M, N)), want)
        _run(dev, lambda c2: c2.fill(bo.ref(), b"\x00"))


def test_fused_bias_dispatch_in_both_command_buffer_kinds(dev, dev_ops):
    M, N, K = 256, 512, 128
    lhs = oracle.generate_random_matrix("bf16", (M, K), 1)
    rhs = oracle.generate_random_matrix("bf16", (K, N), 2)
    bias = oracle.generate_random_matrix("f32", (N,), 4)
    bl, br, bb = _upload(dev, lhs), _upload(dev, rhs), _upload(dev, bias)
    bo = dev.allocate(M * N * 2)
    cb = H.CommandBuffer(dev).begin()
    dev_ops.record_matmul_bias(cb, "bf16", 1, M, N, K, bl.ref(), br.ref(), bb.ref(), bo.ref(),
                               flags=H.MATMUL_RELU | H.MATMUL_TRUNCATE_RESULT)
    cb.end()
    dev.execute(cb)
    dev.wait_idle()
    assert np.array_equal(_download(dev, bo, np.uint16, (M, N)),
                          oracle.matmul_bias("bf16", lhs, rhs, bias, relu=True, truncate=True))


@pytest.mark.parametrize("shape", [(2, 384, 1100), (3, 1, 1024), (40, 1, 512)], ids=str)
def test_attention_dispatch_chains_in_both_command_buffer_kinds(dev, dev_ops, shape):
    """Prefill with the tail split (2 dispatches joined by a barrier), the decode cluster
    kernel and the decode split + merge pair, replayed as stream launches and as graphs."""
    B, M, K2 = shape
    t, D = "f16", 128
    g = np.random.Generator(np.random.PCG64(0xB200 + B + K2))
    mk = lambda *s: oracle.from_f32(t, np.clip(g.standard_normal(s), -4, 4).astype(np.float32))
    q, k, v = mk(B, M, D), mk(B, K2, D), mk(B, K2, D)
    scale = 1.0 / np.sqrt(float(D))
    sel = dev_ops.select_attention(t, B, M, D, K2, D, scale)
    bq, bk, bv = _upload(dev, q), _upload(dev, k), _upload(dev, v)
    bo = dev.allocate(q.nbytes)
    ws = dev.allocate(max(sel["workspace_bytes"], 16))
    cb = H.CommandBuffer(dev).begin()
    dev_ops.record_attention(cb, t, B, M, D, K2, D, scale, bq.ref(), bk.ref(), bv.ref(), bo.ref(),
                             workspace=ws.ref() if sel["workspace_bytes"] else None)
    cb.end()
    want = oracle.to_f32(t, oracle.attention(t, q, k, v, scale, k2_tile=128))
    for _ in range(2):
        dev.execute(cb)
        dev.wait_idle()
        got = oracle.to_f32(t, _download(dev, bo, np.uint16, q.shape))
        assert np.abs(got - want).max() <= 1e-3 * max(np.abs(want).max(), 1.0)
        _run(dev, lambda c2: c2.fill(bo.ref(), b"\x00"))


# ---- command-buffer recording state machine (command_buffer/basic_test.cc) ----------------

def test_recording_state_machine(dev):
    b = dev.allocate(64)
    cb = H.CommandBuffer(dev)
    with pytest.raises(H.Status) as e:
        cb.fill(b.ref(), b"\x00")                    # not recording yet
    assert e.value.name == "FAILED_PRECONDITION"
    cb.begin()
    with pytest.raises(H.Status):
        cb.begin()                                   # already recording
    with pytest.raises(H.Status) as e:
        dev.execute(cb)                              # not ended
    assert e.value.name == "FAILED_PRECONDITION"
    cb.fill(b.ref(), b"\x42")
    cb.end()
    with pytest.raises(H.Status):
        cb.fill(b.ref(), b"\x00")                    # executable: no more recording
    with pytest.raises(H.Status):
        cb.end()
    dev.execute(cb)
    dev.wait_idle()
    assert (_download(dev, b, np.uint8, 64) == 0x42).all()


def test_command_buffer_keeps_its_resources_alive(dev):
    """resource_set semantics (stream_command_buffer.c:528-530): the caller may drop its
    references once the command is recorded."""
    sink = dev.allocate(4096)
    cb = H.CommandBuffer(dev).begin()
    src = _upload(dev, np.arange(4096, dtype=np.uint8) % 251)
    cb.copy(src.ref(), sink.ref())
    cb.end()
    src.release()                                    # only the command buffer holds it now
    dev.execute(cb)
    dev.wait_idle()
    assert np.array_equal(_download(dev, sink, np.uint8, 4096), np.arange(4096, dtype=np.uint8) % 251)


def test_queue_execute_retains_the_command_buffer_until_the_gpu_is_done(dev):
    """The reference keeps a submitted command buffer alive until its signal timepoint
    (cuda_device.c:1168-1196). Here: queue_execute retains it behind an event; the caller may
    release it right after submitting — also in GRAPH mode, where an update_buffer node reads
    the command buffer's captured bytes when the graph runs."""
    n = 1 << 20
    payload = (np.arange(n, dtype=np.uint32) * 2654435761 & 0xFF).astype(np.uint8)
    sink, mid = dev.allocate(n), dev.allocate(n)
    sem = H.Semaphore(dev, 0)
    for round_ in range(3):
        cb = H.CommandBuffer(dev).begin()
        cb.update(bytes(payload ^ round_), mid.ref())
        cb.barrier()
        cb.copy(mid.ref(), sink.ref())
        cb.end()
        dev.execute(cb, signals=[(sem, round_ + 1)])
        cb.release()                                 # the queue holds the only reference now
        del cb
    sem.wait(3)
    assert np.array_equal(_download(dev, sink, np.uint8, n), payload ^ 2)
    dev.wait_idle()


def test_a_submission_that_cannot_be_queued_fails_its_signal_semaphores(dev):
    """A wait on a failed semaphore, or on a value nothing queued will ever signal (no
    wait-before-signal here), is refused AND poisons the signal list, so downstream waiters
    see the failure instead of hanging (cuda_device.c:1187-1191)."""
    bad, never, out1, out2 = (H.Semaphore(dev, 0) for _ in range(4))
    bad.fail(10, "upstream died")                    # ABORTED
    with pytest.raises(H.Status):
        dev.execute(None, waits=[(bad, 1)], signals=[(out1, 1)])
    with pytest.raises(H.Status):
        out1.wait(1, timeout_ns=1_000_000)
    with pytest.raises(H.Status) as e:
        dev.execute(None, waits=[(never, 7)], signals=[(out2, 1)])
    assert e.value.name == "FAILED_PRECONDITION"
    with pytest.raises(H.Status) as e:
        out2.query()
    assert e.value.name == "FAILED_PRECONDITION"


# ---- core/executable_cache_test.cc, executable_test.cc -----------------------------------

def test_executable_cache(dev, dev_ops):
    cache = H.ExecutableCache(dev, "cts")
    assert cache.can_prepare_format("PTXE") and not cache.can_prepare_format("HSACO")
    size = ctypes.c_size_t()
    ptr = H.lib().iree_b200_ops_builtin_executable_data(ctypes.byref(size))
    blob = ctypes.string_at(ptr, size.value)
    exe = cache.prepare(blob)
    builtin = dev_ops.executable
    assert exe.export_count == builtin.export_count > 50
    for name in ("iree_b200_matmul_bf16_f32", "iree_b200_attention_decode_cluster_f16",
                 "iree_b200_matmul_bias_i8_i32_1cta"):
        o = exe.lookup(name)
        assert o == builtin.lookup(name)
        info = exe.export_info(o)
        assert info["kernel_name"] == name and info == builtin.export_info(o)
    assert exe.export_info(exe.lookup("iree_b200_matmul_bf16_f32"))["tma_map_count"] == 3
    assert exe.export_info(exe.lookup("iree_b200_matmul_generic_i8_i32"))["tma_map_count"] == 0
    with pytest.raises(H.Status) as e:
        exe.lookup("no_such_kernel")
    assert e.value.name == "NOT_FOUND"
    with pytest.raises(H.Status):
        exe.export_info(exe.export_count)
    with pytest.raises(H.Status):
        cache.prepare(blob, fmt="HSACO")
    # Untrusted input (native_executable.c:127-133): status, never a crash.
    for bad in (blob[:100], b"\x00" * 512, blob[:64] + bytes(len(blob) - 64)):
        with pytest.raises(H.Status):
            cache.prepare(bad)
    # The second executable is independently usable.
    M = N = K = 32
    lhs = oracle.generate_random_matrix("i8", (M, K), 1)
    rhs = oracle.generate_random_matrix("i8", (K, N), 2)
    bl, br, bo = _upload(dev, lhs), _upload(dev, rhs), dev.allocate(M * N * 4)
    o = exe.lookup("iree_b200_matmul_generic_i8_i32")
    _run(dev, lambda cb: cb.dispatch(exe, o, [1, 1, 1], [M, N, K, 1, 0], [bl.ref(), br.ref(), bo.ref()]))
    assert np.array_equal(_download(dev, bo, np.int32, (M, N)), oracle.matmul("i8", lhs, rhs))


def test_device_queries(dev):
    assert dev.query_i64("hal.device.id", "cuda") == 1 and dev.query_i64("hal.device.id", "hip") == 0
    assert dev.query_i64("hal.executable.format", "cuda-nvptx-fb") == 1
    assert dev.query_i64("hal.executable.format", "vulkan-spirv-fb") == 0
    assert dev.query_i64("cuda.device", "compute_capability_major") == 10
    assert dev.query_i64("cuda.device", "compute_capability_minor") == 0
    assert dev.query_i64("b200.device", "multiprocessor_count") == 148
    assert dev.query_i64("b200.device", "max_shared_memory_per_block_optin") >= 227 * 1024
    with pytest.raises(H.Status) as e:
        dev.query_i64("b200.device", "nope")
    assert e.value.name == "NOT_FOUND"
    assert dev.stream != 0


# ---- core/semaphore_test.cc, queue/semaphore_submission_test.cc --------------------------

def test_semaphores_and_queue_ordering(dev):
    sem = H.Semaphore(dev, 5)
    assert sem.query() == 5
    sem.wait(5)
    sem.wait(3)                                      # already reached
    a = _upload(dev, np.zeros(1 << 22, dtype=np.uint8))
    b = _upload(dev, np.zeros(1 << 22, dtype=np.uint8))
    first = H.CommandBuffer(dev).begin()
    first.fill(a.ref(), b"\x33")
    first.end()
    second = H.CommandBuffer(dev).begin()
    second.copy(a.ref(), b.ref())
    second.end()
    other = H.Semaphore(dev, 0)
    dev.execute(first, signals=[(sem, 6)])
    dev.execute(second, waits=[(sem, 6)], signals=[(sem, 7), (other, 1)])
    other.wait(1)
    assert sem.query() >= 7 and other.query() == 1
    assert (_download(dev, b, np.uint8, 1 << 22) == 0x33).all()
    # Signal-only and wait-only submissions (no command buffer).
    dev.execute(None, signals=[(sem, 8)])
    dev.execute(None, waits=[(sem, 8)])
    sem.wait(8)
    assert sem.query() == 8
    # Nothing queued ever reaches 100: the infinite wait fails instead of hanging, the
    # bounded one times out, and a submission waiting on it is refused.
    with pytest.raises(H.Status) as e:
        sem.wait(100)
    assert e.value.name == "FAILED_PRECONDITION"
    with pytest.raises(H.Status) as e:
        sem.wait(100, timeout_ns=2_000_000)
    assert e.value.name == "DEADLINE_EXCEEDED"
    with pytest.raises(H.Status) as e:
        dev.execute(None, waits=[(sem, 100)])
    assert e.value.name == "FAILED_PRECONDITION"


def test_queue_alloca_and_dealloca_follow_the_timeline(dev):
    """Reference CTS queue/queue_alloca_test.cc: a transient is allocated after its waits,
    usable by the submission that waits on its signal, and released on the timeline."""
    sem = H.Semaphore(dev, 0)
    src = _upload(dev, np.arange(4096, dtype=np.uint8))
    prep = H.CommandBuffer(dev).begin()
    prep.fill(src.ref(0, 1024), b"\x11")
    prep.end()
    dev.execute(prep, signals=[(sem, 1)])
    tmp = dev.queue_alloca(4096, waits=[(sem, 1)], signals=[(sem, 2)])
    assert tmp.byte_length == 4096 and sem.query() == 2
    cb = H.CommandBuffer(dev).begin()
    cb.copy(src.ref(), tmp.ref())
    cb.end()
    dev.execute(cb, waits=[(sem, 2)], signals=[(sem, 3)])
    dev.queue_dealloca(tmp, waits=[(sem, 3)], signals=[(sem, 4)])
    assert sem.query() == 4
    got = _download(dev, tmp, np.uint8, 4096)        # our reference keeps the memory alive
    want = np.arange(4096, dtype=np.uint8)
    want[:1024] = 0x11
    assert np.array_equal(got, want)
    # A failed allocation fails the signal list instead of leaving waiters hanging.
    bad = H.Semaphore(dev, 0)
    with pytest.raises(H.Status):
        dev.queue_alloca(1 << 60, signals=[(bad, 1)])
    with pytest.raises(H.Status):
        bad.wait(1, timeout_ns=1_000_000)
    with pytest.raises(H.Status) as e:
        bad.query()
    assert e.value.name != "DEADLINE_EXCEEDED"
    # Waiting on a failed semaphore refuses the allocation.
    with pytest.raises(H.Status):
        dev.queue_alloca(64, waits=[(bad, 1)])
    dev.wait_idle()


def test_semaphore_host_signal_and_failure(dev):
    """The host side of the timeline (reference CTS core/semaphore_test.cc: NoOp, NormalSignaling,
    Failure, WaitThenFail / FailThenWait; semaphore_thread_test.cc for the cross-thread wait)."""
    import threading
    sem = H.Semaphore(dev, 2)
    sem.signal(3)
    sem.signal(40)                                   # values may skip
    assert sem.query() == 40
    sem.wait(40)
    for stale in (40, 7, 0):                         # each value once, values only grow
        with pytest.raises(H.Status) as e:
            sem.signal(stale)
        assert e.value.name == "INVALID_ARGUMENT"
    assert sem.query() == 40
    # A submission issued after the host signal sees it; device signals continue from there.
    buf = _upload(dev, np.zeros(4096, dtype=np.uint8))
    cb = H.CommandBuffer(dev).begin()
    cb.fill(buf.ref(), b"\x5a")
    cb.end()
    dev.execute(cb, waits=[(sem, 40)], signals=[(sem, 41)])
    sem.wait(41)
    assert (_download(dev, buf, np.uint8, 4096) == 0x5A).all()
    # Another thread's bounded wait is released by a host signal.
    result = []

    def waiter():
        try:
            sem.wait(50, timeout_ns=5_000_000_000)
            result.append("ok")
        except H.Status as err:          # pragma: no cover - reported through the assert below
            result.append(err.name)

    th = threading.Thread(target=waiter)
    th.start()
    sem.signal(50)
    th.join(10)
    assert result == ["ok"]
    # Failure: the first one wins and poisons query / wait / signal / submissions.
    ABORTED, DATA_LOSS = 10, 15
    sem.fail(ABORTED, "upstream producer died")
    sem.fail(DATA_LOSS, "second failure is ignored")
    for call in (sem.query, lambda: sem.wait(50), lambda: sem.wait(60, timeout_ns=1_000_000),
                 lambda: sem.signal(60), lambda: dev.execute(None, waits=[(sem, 60)])):
        with pytest.raises(H.Status) as e:
            call()
        assert e.value.name == "ABORTED" and "upstream producer died" in e.value.message
    with pytest.raises(H.Status):
        H.Semaphore(dev, 0).fail(0)                  # OK is not a failure
    dev.wait_idle()


def test_many_small_submissions(dev, dev_ops):
    """command_buffer/stress_test.cc in miniature: 200 submissions of a two-command buffer,
    each ordered behind the previous one through one timeline."""
    n = 4096
    acc = _upload(dev, np.zeros(n, dtype=np.int32))
    ones = _upload(dev, np.ones((n, 1), dtype=np.int8))
    one = _upload(dev, np.ones((1, 1), dtype=np.int8))
    exe, ordinal = _generic_i8(dev_ops)
    # acc[n,1] += ones[n,1] . one[1,1]
    cb = H.CommandBuffer(dev).begin()
    cb.dispatch(exe, ordinal, [1, n // 64, 1], [n, 1, 1, 1, 1], [ones.ref(), one.ref(), acc.ref()])
    cb.barrier()
    cb.fill(ones.ref(), b"\x01")
    cb.end()
    sem = H.Semaphore(dev, 0)
    for i in range(200):
        dev.execute(cb, waits=[(sem, i)] if i else (), signals=[(sem, i + 1)])
    sem.wait(200)
    assert (_download(dev, acc, np.int32, n) == 200).all()


# ---- collectives on a single-rank channel --------------------------------------------------

def test_single_rank_channel_collectives(dev):
    """iree_hal_channel_t over NCCL with count == 1 (what a 1-GPU run of a sharded module
    sees): every collective degenerates to a copy. In graph command buffers each batch of
    collectives is stream-captured into a child-graph node (the reference leaves this
    UNIMPLEMENTED, graph_command_buffer.c:291-327)."""
    try:
        ch = H.Channel(dev, rank=0, count=1, id_bytes=H.Channel.create_id())
    except H.Status as e:
        if e.name == "UNAVAILABLE":
            pytest.skip("NCCL not loadable: " + e.message)
        raise
    assert ch.rank_and_count() == (0, 1)
    src = _upload(dev, np.arange(1024, dtype=np.float32))
    dst = _upload(dev, np.zeros(1024, dtype=np.float32))
    cb = H.CommandBuffer(dev).begin()
    op = H.collective_op(H.ALL_REDUCE, H.REDUCTION_SUM, "f32")
    cb.collective(ch, op, 0, src.ref(), dst.ref(), 1024)
    cb.barrier()
    gather = _upload(dev, np.zeros(1024, dtype=np.float32))
    cb.collective(ch, H.collective_op(H.ALL_GATHER, H.REDUCTION_NONE, "f32"), 0, dst.ref(),
                  gather.ref(), 1024)
    cb.barrier()
    cb.fill(src.ref(), b"\x00")                     # ordered behind both collectives
    cb.end()
    ramp = np.arange(1024, dtype=np.float32)
    for _ in range(2):                               # replayable
        dev.h2d(ramp.ctypes.data, src, 4096)
        _run(dev, lambda c2: (c2.fill(dst.ref(), b"\x00"), c2.fill(gather.ref(), b"\x00")))
        dev.execute(cb)
        dev.wait_idle()
        assert np.array_equal(_download(dev, gather, np.float32, 1024),
                              np.arange(1024, dtype=np.float32))
        assert (_download(dev, src, np.float32, 1024) == 0).all()
    # Stream mode issues the batches on every submission; graph mode captured them once.
    assert ch.collective_counts()["nccl"] == (2 if dev._mode_name == "graph" else 4)
    bad = H.CommandBuffer(dev).begin()
    with pytest.raises(H.Status) as e:   # recv binding too small for count * elements
        bad.collective(ch, op, 0, src.ref(), dst.ref(0, 512), 1024)
    assert e.value.name == "OUT_OF_RANGE"
    with pytest.raises(H.Status):        # element type 200 does not exist
        bad.collective(ch, H.collective_op(H.ALL_REDUCE, H.REDUCTION_SUM, 200), 0, src.ref(),
                       dst.ref(), 16)

