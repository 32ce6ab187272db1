"""Launch-layer checks that only make sense under the mock driver (tests/mock_cuda/mock_cuda.cc):
what reaches cuLaunchKernel / cuTensorMapEncodeTiled / the graph API for the ops layer's
dispatches. Run by tests/test_hal_mock_cpu.py in a subprocess whose libcuda.so.1 is the mock;
not collected by the normal test run (the file name does not match test_*.py)."""

import ctypes

import numpy as np
import pytest

from iree_b200 import hal as H

pytestmark = pytest.mark.gpu   # "needs a device": here the device is the mock

MODES = {"stream": H.MODE_STREAM, "graph": H.MODE_GRAPH}


def _mock():
    lib = ctypes.CDLL("libcuda.so.1")
    assert hasattr(lib, "mock_cuda_take_launches"), "the real driver is loaded, not the mock"
    for fn in (lib.mock_cuda_take_launches, lib.mock_cuda_take_tensor_maps):
        fn.restype = ctypes.c_size_t
        fn.argtypes = [ctypes.c_char_p, ctypes.c_size_t]
    return lib


def _take(fn):
    buf = ctypes.create_string_buffer(1 << 20)
    n = fn(buf, len(buf))
    return [line.split() for line in buf.raw[:n].decode().splitlines()]


def launches():
    return [(l[0], [int(x) for x in l[1:4]], [int(x) for x in l[4:7]], int(l[7]))
            for l in _take(_mock().mock_cuda_take_launches)]


def launch_args():
    """(kernel name, [first four parameter slots]) per launch (matmul dispatches only)."""
    return [(l[0], [int(x) for x in l[8:12]]) for l in _take(_mock().mock_cuda_take_launches)]


def tensor_maps():
    out = []
    for l in _take(_mock().mock_cuda_take_tensor_maps):
        v = [int(x) for x in l]
        out.append({"type": v[0], "rank": v[1], "swizzle": v[2], "base": v[3],
                    "dims": v[4:9], "strides": v[9:14], "box": v[14:19]})
    return out


@pytest.fixture(scope="module", params=list(MODES), ids=list(MODES))
def dev(request):
    return H.Device(0, mode=MODES[request.param])


@pytest.fixture(scope="module")
def dev_ops(dev):
    return H.Ops(dev)


def test_mock_device_is_a_b200(dev):
    assert dev.query_i64("cuda.device", "compute_capability_major") == 10
    assert dev.query_i64("b200.device", "multiprocessor_count") == 148


def test_matmul_dispatch_reaches_the_driver_as_recorded(dev, dev_ops):
    """One 2-CTA bf16 dispatch: persistent grid, block, opt-in shared memory, and the three
    tensor maps the `__tma` recipe describes (read from the cubin's ELF symbol table),
    evaluated from (binding pointers, constants)."""
    M, N, K = 8192, 4096, 2048
    a, b, c = dev.allocate(M * K * 2), dev.allocate(K * N * 2), dev.allocate(M * N * 4)
    launches(), tensor_maps()   # drain
    cb = H.CommandBuffer(dev).begin()
    dev_ops.record_matmul(cb, "bf16", 1, M, N, K, a.ref(), b.ref(), c.ref())
    cb.end()
    maps = tensor_maps()        # encoded at record time
    assert launches() == []     # nothing launches before queue_execute
    for _ in range(2):
        dev.execute(cb)
    dev.wait_idle()
    info = dev_ops.executable.export_info(dev_ops.executable.lookup("iree_b200_matmul_bf16_f32"))
    want = ("iree_b200_matmul_bf16_f32", [148, 1, 1], [256, 1, 1], info["block_shared_memory_size"])
    assert launches() == [want, want]
    assert info["block_shared_memory_size"] > 48 * 1024 and info["tma_map_count"] == 3
    ta, tb, tc = maps
    assert ta["base"] == a.device_pointer and tb["base"] == b.device_pointer and tc["base"] == c.device_pointer
    # A [batch, M, K] bf16: innermost K, 128-byte (64-element) x 128-row boxes, 128B swizzle.
    assert (ta["type"], ta["rank"], ta["swizzle"]) == (9, 3, 3)
    assert ta["dims"][:3] == [K, M, 1] and ta["strides"][:2] == [K * 2, M * K * 2]
    assert ta["box"][:3] == [64, 128, 1]
    # B [batch, K, N] (MN-major operand): innermost N in 64-element boxes of 64 k rows.
    assert tb["dims"][:3] == [N, K, 1] and tb["strides"][:2] == [N * 2, K * N * 2]
    assert tb["box"][:3] == [64, 64, 1]
    # C [batch, M, N] f32: 32 columns (128 bytes) x 128 rows per store.
    assert (tc["type"], tc["dims"][:3], tc["box"][:3]) == (7, [N, M, 1], [32, 128, 1])
    assert tc["strides"][:2] == [N * 4, M * N * 4]


def test_sub_range_bindings_move_the_tensor_map_base(dev, dev_ops):
    M = N = K = 512
    big = dev.allocate(4 << 20)
    cb = H.CommandBuffer(dev).begin()
    tensor_maps()
    dev_ops.record_matmul(cb, "i8", 1, M, N, K, big.ref(1024, M * K), big.ref(1 << 20, K * N),
                          big.ref(2 << 20, M * N * 4), flags=H.MATMUL_TRANSPOSE_RHS)
    ta, tb, tc = tensor_maps()
    assert [t["base"] - big.device_pointer for t in (ta, tb, tc)] == [1024, 1 << 20, 2 << 20]
    assert ta["type"] == 0 and ta["box"][:2] == [128, 128]      # bytes: 128 per row
    assert tb["dims"][:2] == [K, N]                             # [N, K]: K-major rhs
    assert tc["type"] == 3                                      # i32 result


def test_device_map_twins_reach_the_driver_without_tensor_maps(dev, dev_ops):
    """FLAG_DEVICE_MAPS: the `<name>_dtm` export is launched with the ABI parameters only —
    nothing is encoded on the host (no cuTensorMapEncodeTiled call), same grid, block and
    shared memory as the host-fed twin. This is what the stock IREE CUDA driver does."""
    M, N, K = 4096, 4096, 1024
    a, b, c = dev.allocate(M * K * 2), dev.allocate(K * N * 2), dev.allocate(M * N * 4)
    bias = dev.allocate(N * 4)
    tensor_maps(), launches()
    for flags, record in ((0, None), (H.MATMUL_RELU | H.MATMUL_TRUNCATE_RESULT, "bias")):
        got = {}
        for device_maps in (False, True):
            f = flags | (H.MATMUL_DEVICE_MAPS if device_maps else 0)
            cb = H.CommandBuffer(dev).begin()
            if record == "bias":
                dev_ops.record_matmul_bias(cb, "bf16", 1, M, N, K, a.ref(), b.ref(), bias.ref(), c.ref(0, M * N * 2),
                                           flags=f)
            else:
                dev_ops.record_matmul(cb, "bf16", 1, M, N, K, a.ref(), b.ref(), c.ref(), flags=f)
            cb.end()
            maps = tensor_maps()
            dev.execute(cb)
            dev.wait_idle()
            (launch,) = launches()
            got[device_maps] = (launch, len(maps))
        (host_fed, n_host), (twin, n_twin) = got[False], got[True]
        assert n_host == 3 and n_twin == 0
        assert twin[0] == host_fed[0] + "_dtm" and twin[1:] == host_fed[1:]


@pytest.mark.parametrize("t,M0,N0", [("bf16", 16, 16), ("i8", 8, 32), ("f16", 128, 64)])
def test_mmt4d_tile_direct_maps_walk_the_tiles(dev, dev_ops, t, M0, N0):
    """linalg.mmt4d with 16-byte K0 rows and tile rows in multiples of 8: ONE tcgen05 dispatch
    whose rank-4 tensor maps read the tiled operands as UMMA core matrices (8 rows x 16 bytes =
    128 contiguous bytes, unswizzled) and write result tiles with a swizzle as wide as the
    tile row; outside the guard the call stays on the SIMT / relayout routes."""
    es = 1 if t == "i8" else 2
    K0 = 16 // es
    M, N, K = 1024, 2048, 512
    M1, N1, K1 = M // M0, N // N0, K // K0
    a, b, c = dev.allocate(M * K * es), dev.allocate(N * K * es), dev.allocate(M * N * 4)
    tensor_maps(), launches()
    cb = H.CommandBuffer(dev).begin()
    dev_ops.record_mmt4d(cb, t, M1, N1, K1, M0, N0, K0, a.ref(), b.ref(), c.ref())
    cb.end()
    ta, tb, tc = tensor_maps()
    dev.execute(cb)
    dev.wait_idle()
    acc = "i32" if t == "i8" else "f32"
    (name, grid, block, smem), = launches()
    assert name == f"iree_b200_mmt4d_{t}_{acc}_1cta" and block == [256, 1, 1] and grid == [64, 1, 1]
    inner = 128 // es          # 8 rows x K0 elements
    for tm, R0, R1, rows in ((ta, M0, M1, 128), (tb, N0, N1, 256)):
        assert (tm["rank"], tm["swizzle"]) == (4, 0)
        assert tm["dims"][:4] == [inner, K1, R0 // 8, R1]
        # byte strides of (k1, 8-row group, r1) in the tiled tensor [R1][K1][R0 / 8][8 x K0]
        assert tm["strides"][:3] == [R0 * 16, 128, K1 * R0 * 16]
        assert tm["box"][:4] == [inner, 8, R0 // 8, rows // R0]
    cols = min(N0, 32)
    assert (tc["rank"], tc["swizzle"], tc["type"]) == (4, {32: 3, 16: 2, 8: 1}[cols], 3 if t == "i8" else 7)
    assert tc["dims"][:4] == [N0, N1, M0, M1]
    assert tc["strides"][:3] == [M0 * N0 * 4, N0 * 4, N1 * M0 * N0 * 4]
    assert tc["box"][:4] == [cols, 32 // cols, M0, 128 // M0]
    # Outside the guard (tile rows not a multiple of 8; a K0 row that is not 16 bytes): no maps.
    for m0, n0, k0 in ((4, 16, K0), (16, 16, K0 * 2)):
        cb = H.CommandBuffer(dev).begin()
        dev_ops.record_mmt4d(cb, t, M // m0, N // n0, K // k0, m0, n0, k0, a.ref(), b.ref(), c.ref())
        cb.end()
        assert tensor_maps() == []


def test_fused_bias_dispatch_binds_the_result_fourth(dev, dev_ops):
    M, N, K = 1024, 1024, 256
    a, b, c = dev.allocate(M * K * 2), dev.allocate(K * N * 2), dev.allocate(M * N * 2)
    bias = dev.allocate(N * 4 + 4)
    cb = H.CommandBuffer(dev).begin()
    tensor_maps(), launches()
    dev_ops.record_matmul_bias(cb, "bf16", 1, M, N, K, a.ref(), b.ref(), bias.ref(4, N * 4), c.ref(),
                               flags=H.MATMUL_RELU | H.MATMUL_TRUNCATE_RESULT)
    cb.end()
    maps = tensor_maps()
    assert len(maps) == 3 and maps[2]["base"] == c.device_pointer
    assert maps[2]["type"] == 9 and maps[2]["box"][:2] == [64, 128]   # bf16 result, 64 columns
    dev.execute(cb)
    # 16 tiles of 256x256 would leave most SM pairs idle: the 1-CTA family takes it.
    assert [l[0] for l in launches()] == ["iree_b200_matmul_bias_bf16_bf16_1cta"]


def test_attention_chains_keep_their_order(dev, dev_ops):
    """Prefill with a tail split = split kernel, barrier, merge kernel; decode with few rows =
    split + merge; decode with many rows = one cluster launch of 8 CTAs per row."""
    def run(B, M, K2):
        es = 2
        q, o = dev.allocate(B * M * 128 * es), dev.allocate(B * M * 128 * es)
        k, v = dev.allocate(B * K2 * 128 * es), dev.allocate(B * K2 * 128 * es)
        sel = dev_ops.select_attention("f16", B, M, 128, K2, 128, 0.088)
        ws = dev.allocate(max(16, sel["workspace_bytes"]))
        launches()
        cb = H.CommandBuffer(dev).begin()
        dev_ops.record_attention(cb, "f16", B, M, 128, K2, 128, 0.088, q.ref(), k.ref(), v.ref(),
                                 o.ref(), workspace=ws.ref() if sel["workspace_bytes"] else None)
        cb.end()
        dev.execute(cb)
        dev.wait_idle()
        return launches(), sel

    got, sel = run(32, 4096, 4096)
    assert [l[0] for l in got] == ["iree_b200_attention_prefill_split_f16",
                                   "iree_b200_attention_prefill_merge_f16"]
    assert got[0][1] == [444 + 68 * 2, 1, 1] and got[1][1] == [32, 68, 1]
    got, _ = run(4, 1, 8192)
    assert [l[0] for l in got] == ["iree_b200_attention_decode_split_f16",
                                   "iree_b200_attention_decode_merge_f16"]
    got, _ = run(32, 1, 8192)
    assert [(l[0], l[1], l[2]) for l in got] == [("iree_b200_attention_decode_cluster_f16",
                                                  [8, 1, 32], [288, 1, 1])]


def test_pipelined_host_call_issues_one_dispatch_per_block(dev, dev_ops):
    """iree_b200_ops_matmul_host above the pipeline threshold: panels go up, one dispatch per
    (row panel, column panel) block, in diagonal order; the mock does not compute, so only
    the launch sequence is checked here (results: tests/test_matmul_gpu.py on a GPU)."""
    M, N, K = 1024, 768, 64
    lhs = np.zeros((M, K), dtype=np.uint16)
    rhs = np.zeros((K, N), dtype=np.uint16)
    out = np.zeros((M, N), dtype=np.float32)
    dev_ops.set_host_pipeline(0, 256, 256)
    launches()
    try:
        dev_ops.matmul_host("bf16", 1, M, N, K, lhs.ctypes.data, rhs.ctypes.data, out.ctypes.data)
    finally:
        dev_ops.set_host_pipeline(16 << 20, 2048, 2048)
    got = launches()
    assert len(got) == (M // 256) * (N // 256)
    assert {l[0] for l in got} == {"iree_b200_matmul_bf16_f32_1cta"}
    assert all(l[1] == [2, 1, 1] for l in got)      # 256x256 block = two 128x256 tiles


@pytest.mark.parametrize("case", [(1, 1000, 272, 528, False, False), (2, 513, 64, 784, True, True)],
                         ids=str)
def test_pipelined_bias_call_hands_each_block_its_bias_columns(dev, dev_ops, case):
    """iree_b200_ops_matmul_bias_host through the pipeline: the bias goes up once (the mock's
    device memory is host memory, so the staged copy is read back here), each block's
    dispatch is the fused export with binding 2 at bias + 4 * (first column of the block),
    and lhs / result bindings follow the row panel / packed-block layout."""
    batch, M, K, N, nt, truncate = case
    block, out_es = 256, 2 if truncate else 4
    lhs = np.zeros((batch, M, K), dtype=np.uint16)
    rhs = np.zeros((batch, N, K) if nt else (batch, K, N), dtype=np.uint16)
    bias = np.arange(N, dtype=np.float32) + 0.5
    guard = 64
    raw = np.full(batch * M * N * out_es + 2 * guard, 0xA5, dtype=np.uint8)
    flags = (H.MATMUL_RELU | (H.MATMUL_TRANSPOSE_RHS if nt else 0)
             | (H.MATMUL_TRUNCATE_RESULT if truncate else 0))
    dev_ops.set_host_pipeline(0, block, block)
    launches()
    try:
        dev_ops.matmul_bias_host("bf16", batch, M, N, K, lhs.ctypes.data, rhs.ctypes.data,
                                 bias.ctypes.data, raw.ctypes.data + guard, flags)
    finally:
        dev_ops.set_host_pipeline(16 << 20, 2048, 2048)
    got = launch_args()
    rows, cols = -(-M // block), -(-N // block)
    assert len(got) == batch * rows * cols
    suffix = ("_nt" if nt else "") + ("_bf16_bf16" if truncate else "_bf16_f32")
    assert {g[0].replace("_1cta", "") for g in got} == {"iree_b200_matmul_bias" + suffix}
    assert (raw[:guard] == 0xA5).all() and (raw[-guard:] == 0xA5).all()

    lhs0 = min(g[1][0] for g in got)
    bias0 = min(g[1][2] for g in got)
    out0 = min(g[1][3] for g in got)
    staged = np.ctypeslib.as_array(ctypes.cast(bias0, ctypes.POINTER(ctypes.c_float)), (N,))
    assert np.array_equal(staged, bias)
    # The bias sits behind the (1 KiB-rounded) result staging.
    assert bias0 == out0 + -(-batch * M * N * out_es // 1024) * 1024
    seen = set()
    for _, (a, _, bptr, c) in got:
        j = (bptr - bias0) // (4 * block)
        assert bptr == bias0 + 4 * block * j and 0 <= j < cols
        b, rem = divmod(a - lhs0, M * K * 2)
        i = rem // (block * K * 2)
        assert rem == i * block * K * 2 and 0 <= b < batch and 0 <= i < rows
        mb = min(block, M - i * block)
        assert c == out0 + ((b * M + i * block) * N + mb * j * block) * out_es
        seen.add((b, i, j))
    assert len(seen) == len(got)                       # every block exactly once


def test_pending_executable_loads_on_first_use_only(dev, dev_ops):
    M, N, K = 64, 64, 64
    z = lambda n: np.zeros(n, dtype=np.uint8)
    lhs, ls, rhs, rs = z(M * K // 2), z(M * K // 32), z(N * K // 2), z(N * K // 32)
    out = np.zeros((M, N), dtype=np.float32)
    launches()
    dev_ops.matmul_mxfp4_host(M, N, K, lhs.ctypes.data, ls.ctypes.data, rhs.ctypes.data,
                              rs.ctypes.data, out.ctypes.data)
    assert launches() == [("iree_b200_matmul_generic_mxfp4_f32", [1, 1, 1], [256, 1, 1], 0)]


@pytest.mark.parametrize("case", [
    # batch, M, K, N, accumulate, nt, truncate, block
    (1, 1000, 272, 528, True, False, False, 256),
    (1, 528, 144, 1008, False, True, False, 256),
    (3, 300, 528, 272, False, False, False, 256),
    (2, 513, 64, 768, False, False, True, 256),
    (1, 1536, 144, 1280, False, False, False, 512),
], ids=str)
def test_pipelined_matmul_staging_stays_in_bounds(dev, dev_ops, case):
    """Ragged panels, batches, transpose_rhs, accumulate and truncation through the three-stream
    host pipeline. The mock's copies are real memcpys between exactly-sized allocations, so
    under tools/sanitize_host.sh (ASAN) any staging offset error is an abort; here the
    launch count and the untouched guard bytes around the result are asserted."""
    batch, M, K, N, accumulate, nt, truncate, block = case
    lhs = np.zeros((batch, M, K), dtype=np.uint16)
    rhs = np.zeros((batch, N, K) if nt else (batch, K, N), dtype=np.uint16)
    guard = 64
    raw = np.full(batch * M * N * (2 if truncate else 4) + 2 * guard, 0xA5, dtype=np.uint8)
    out_ptr = raw.ctypes.data + guard
    flags = ((H.MATMUL_ACCUMULATE if accumulate else 0) | (H.MATMUL_TRANSPOSE_RHS if nt else 0)
             | (H.MATMUL_TRUNCATE_RESULT if truncate else 0))
    dev_ops.set_host_pipeline(0, block, block)
    launches()
    try:
        dev_ops.matmul_host("bf16", batch, M, N, K, lhs.ctypes.data, rhs.ctypes.data, out_ptr, flags)
    finally:
        dev_ops.set_host_pipeline(16 << 20, 2048, 2048)
    assert len(launches()) == batch * -(-M // block) * -(-N // block)
    assert (raw[:guard] == 0xA5).all() and (raw[-guard:] == 0xA5).all()


@pytest.mark.parametrize("case", [("f16", 5, 130, 128, 300, 128, None),
                                  ("bf16", 3, 512, 128, 1025, 128, None),
                                  ("f32", 4, 33, 24, 50, 40, "bool"),
                                  ("f32", 3, 20, 16, 70, 8, "f32")], ids=str)
def test_sliced_attention_staging_stays_in_bounds(dev, dev_ops, case):
    t, B, M, K1, K2, N, mask_kind = case
    es = 4 if t == "f32" else 2
    q = np.zeros(B * M * K1 * es, dtype=np.uint8)
    k = np.zeros(B * K2 * K1 * es, dtype=np.uint8)
    v = np.zeros(B * K2 * N * es, dtype=np.uint8)
    mask = None if mask_kind is None else np.zeros(B * M * K2 * (1 if mask_kind == "bool" else 4),
                                                   dtype=np.uint8)
    guard = 64
    raw = np.full(B * M * N * es + 2 * guard, 0xA5, dtype=np.uint8)
    dev_ops.set_host_pipeline(0, 2048, 2048)
    launches()
    try:
        dev_ops.attention_host(t, B, M, K1, K2, N, 0.1, q.ctypes.data, k.ctypes.data, v.ctypes.data,
                               raw.ctypes.data + guard,
                               mask_ptr=0 if mask is None else mask.ctypes.data,
                               mask_kind={None: 0, "bool": 1, "f32": 2}[mask_kind])
    finally:
        dev_ops.set_host_pipeline(16 << 20, 2048, 2048)
    per_slice = len(dev_ops.select_attention(t, 1, M, K1, K2, N, 0.1,
                                             mask_kind={None: 0, "bool": 1, "f32": 2}[mask_kind])["dispatches"])
    assert len(launches()) == B * per_slice
    assert (raw[:guard] == 0xA5).all() and (raw[-guard:] == 0xA5).all()


# ---- with MOCK_CUDA_KERNELS: the SIMT kernels run (their own source compiled for the host) ----

def _kernels_run():
    import os
    return bool(os.environ.get("MOCK_CUDA_KERNELS"))


needs_kernels = pytest.mark.skipif(not _kernels_run(), reason="MOCK_CUDA_KERNELS not set")


@needs_kernels
@pytest.mark.parametrize("t", ["i8", "f16", "bf16", "f8e4m3fn", "f8e5m2"])
@pytest.mark.parametrize("case", [(1, 70, 33, 50, False, False), (2, 17, 40, 65, True, True)], ids=str)
def test_generic_matmul_end_to_end_on_the_host(dev, dev_ops, t, case):
    """Host call -> staging -> record -> dispatch parameter packing -> (mock) launch -> the
    generic kernel's own code on OS threads -> download, against the oracle, bit for bit: the
    whole ABI path of a dispatch without a GPU."""
    import oracle
    batch, M, K, N, accumulate, nt = case
    lhs = oracle.generate_random_matrix(t, (batch, M, K), 1)
    rhs = oracle.generate_random_matrix(t, (batch, N, K) if nt else (batch, K, N), 2)
    acc_t = "i32" if t == "i8" else "f32"
    acc = oracle.generate_random_matrix(acc_t, (batch, M, N), 3) if accumulate else None
    out = acc.copy() if accumulate else np.full((batch, M, N), 77, dtype=np.int32 if t == "i8" else np.float32)
    flags = (H.MATMUL_FORCE_GENERIC | (H.MATMUL_ACCUMULATE if accumulate else 0)
             | (H.MATMUL_TRANSPOSE_RHS if nt else 0))
    dev_ops.matmul_host(t, batch, M, N, K, lhs.ctypes.data, rhs.ctypes.data, out.ctypes.data, flags)
    assert np.array_equal(out, oracle.matmul(t, lhs, rhs, acc=acc, transpose_rhs=nt))


@needs_kernels
@pytest.mark.parametrize("t", ["i8", "bf16"])
def test_generic_bias_relu_truncate_end_to_end_on_the_host(dev, dev_ops, t):
    import oracle
    M, K, N = 37, 24, 70
    lhs = oracle.generate_random_matrix(t, (1, M, K), 1)
    rhs = oracle.generate_random_matrix(t, (1, K, N), 2)
    bias = oracle.generate_random_matrix("i32" if t == "i8" else "f32", (N,), 4)
    truncate = t != "i8"
    out = np.zeros((1, M, N), dtype=np.uint16 if truncate else np.int32)
    dev_ops.matmul_bias_host(t, 1, M, N, K, lhs.ctypes.data, rhs.ctypes.data, bias.ctypes.data,
                             out.ctypes.data, H.MATMUL_FORCE_GENERIC | H.MATMUL_RELU
                             | (H.MATMUL_TRUNCATE_RESULT if truncate else 0))
    assert np.array_equal(out, oracle.matmul_bias(t, lhs, rhs, bias, relu=True, truncate=truncate))


@needs_kernels
@pytest.mark.parametrize("accumulate", [False, True], ids=["", "acc"])
def test_mxfp4_end_to_end_on_the_host(dev, dev_ops, accumulate):
    """The MXFP4 entry points (second executable, five bindings) down to the kernel's own code."""
    import oracle
    M, K, N = 70, 96, 50
    lhs = oracle.generate_random_matrix("u8", (M, K // 2), 1)
    ls = oracle.generate_random_matrix("f8e8m0fnu", (M, K // 32), 2)
    rhs = oracle.generate_random_matrix("u8", (N, K // 2), 3)
    rs = oracle.generate_random_matrix("f8e8m0fnu", (N, K // 32), 4)
    acc = oracle.generate_random_matrix("f32", (M, N), 5) if accumulate else None
    out = acc.copy() if accumulate else np.full((M, N), 77, dtype=np.float32)
    dev_ops.matmul_mxfp4_host(M, N, K, lhs.ctypes.data, ls.ctypes.data, rhs.ctypes.data,
                              rs.ctypes.data, out.ctypes.data,
                              flags=H.MATMUL_ACCUMULATE if accumulate else 0)
    assert np.array_equal(out, oracle.matmul_mxfp4(lhs, ls, rhs, rs, acc=acc))
    out = np.full((M, N), 5, dtype=np.float32)
    dev_ops.matmul_mxfp4_host(M, N, 0, 0, 0, 0, 0, out.ctypes.data)      # K == 0
    assert (out == 0).all()


@needs_kernels
@pytest.mark.parametrize("t", ["bf16", "i8"])
@pytest.mark.parametrize("case", [
    # batch, M, K, N, accumulate, nt, truncate, bias, relu
    (1, 300, 16, 528, False, False, False, False, False),
    (2, 260, 32, 272, True, True, False, False, False),
    (1, 520, 16, 272, False, False, True, False, False),
    (1, 300, 16, 528, False, False, False, True, True),
    (2, 260, 32, 272, True, True, False, True, False),
    (1, 264, 16, 528, False, True, True, True, True),
], ids=str)
def test_pipelined_host_call_end_to_end_on_the_host(dev, dev_ops, t, case):
    """The three-stream pipeline's whole data path (row / column panel uploads, accumulator
    and bias staging, packed result blocks, 2-D downloads) with the generic kernel standing
    in for the tcgen05 one (threshold 0 lets the forced family through), bit for bit
    against the oracle."""
    import oracle
    batch, M, K, N, accumulate, nt, truncate, with_bias, relu = case
    if truncate and t == "i8":
        pytest.skip("TRUNCATE_RESULT is f16 / bf16 only")
    acc_t = "i32" if t == "i8" else "f32"
    lhs = oracle.generate_random_matrix(t, (batch, M, K), 1)
    rhs = oracle.generate_random_matrix(t, (batch, N, K) if nt else (batch, K, N), 2)
    acc = oracle.generate_random_matrix(acc_t, (batch, M, N), 3) if accumulate else None
    bias = oracle.generate_random_matrix(acc_t, (N,), 4) if with_bias else None
    if truncate:
        out = np.full((batch, M, N), 0x7777, dtype=np.uint16)
    else:
        out = acc.copy() if accumulate else np.full((batch, M, N), 77, dtype=np.dtype(
            np.int32 if t == "i8" else np.float32))
    flags = (H.MATMUL_FORCE_GENERIC | (H.MATMUL_ACCUMULATE if accumulate else 0)
             | (H.MATMUL_TRANSPOSE_RHS if nt else 0) | (H.MATMUL_TRUNCATE_RESULT if truncate else 0)
             | (H.MATMUL_RELU if relu else 0))
    dev_ops.set_host_pipeline(0, 256, 256)
    launches()
    try:
        if with_bias:
            dev_ops.matmul_bias_host(t, batch, M, N, K, lhs.ctypes.data, rhs.ctypes.data,
                                     bias.ctypes.data, out.ctypes.data, flags)
        else:
            dev_ops.matmul_host(t, batch, M, N, K, lhs.ctypes.data, rhs.ctypes.data,
                                out.ctypes.data, flags)
    finally:
        dev_ops.set_host_pipeline(16 << 20, 2048, 2048)
    assert len(launches()) == batch * -(-M // 256) * -(-N // 256)     # it was the pipeline
    if with_bias:
        want = oracle.matmul_bias(t, lhs, rhs, bias, acc=acc, transpose_rhs=nt, relu=relu,
                                  truncate=truncate)
    else:
        want = oracle.matmul(t, lhs, rhs, acc=acc, transpose_rhs=nt)
        if truncate:
            want = oracle.from_f32(t, want)
    assert np.array_equal(out, want)


@needs_kernels
@pytest.mark.parametrize("t", ["i8", "f16", "bf16", "f32"])
@pytest.mark.parametrize("case", [
    # M, K, N (logical, before packing), M0, N0, K0, accumulate
    (37, 50, 29, 16, 16, 2, False),      # the AVX-512-BF16 / VNNI tile shape
    (70, 33, 130, 8, 8, 1, True),
    (64, 64, 64, 16, 16, 16, False),     # an MMA-intrinsic-shaped tile
    (100, 40, 33, 8, 32, 4, True),
    (33, 17, 40, 1, 1, 1, False),        # degenerate tiles: plain [M,K] x [N,K]
    (20, 24, 90, 7, 5, 3, True),         # tiles that divide nothing
], ids=str)
def test_mmt4d_end_to_end_on_the_host(dev, dev_ops, t, case):
    """linalg.mmt4d through the host call on packed operands, bit for bit against the oracle's
    restatement and — when oracle/_ref is built — the reference's own iree_uk_mmt4d; unpacking
    the result gives the plain matmul of the unpacked operands."""
    import oracle
    M, K, N, M0, N0, K0, accumulate = case
    lhs = oracle.generate_random_matrix(t, (M, K), 1)
    rhs = oracle.generate_random_matrix(t, (K, N), 2)
    l4 = oracle.pack(lhs, M0, K0)
    r4 = oracle.pack(np.ascontiguousarray(rhs.T), N0, K0)
    M1, K1, N1 = l4.shape[0], l4.shape[1], r4.shape[0]
    acc_t = "i32" if t == "i8" else "f32"
    acc = oracle.generate_random_matrix(acc_t, (M1, N1, M0, N0), 3) if accumulate else None
    out = acc.copy() if accumulate else np.full((M1, N1, M0, N0), 77, dtype=np.dtype(
        np.int32 if t == "i8" else np.float32))
    launches()
    dev_ops.mmt4d_host(t, M1, N1, K1, M0, N0, K0, l4.ctypes.data, r4.ctypes.data, out.ctypes.data,
                       flags=H.MATMUL_ACCUMULATE if accumulate else 0)
    got = launches()
    assert [(g[0], g[1]) for g in got] == [
        ("iree_b200_mmt4d_generic_%s_%s" % (t, acc_t), [-(-N1 * N0 // 64), -(-M1 * M0 // 64), 1])]
    assert np.array_equal(out, oracle.mmt4d(t, l4, r4, acc=acc))
    if oracle.ref_available():
        assert np.array_equal(out, oracle.ref_mmt4d(t, l4, r4, acc=acc))
    if not accumulate:
        want_mm = (lhs.astype(np.float64) @ rhs.astype(np.float64)).astype(np.float32) if t == "f32" \
            else oracle.matmul(t, lhs[None], rhs[None])[0]
        assert np.array_equal(oracle.unpack(out, M, N), want_mm)


@needs_kernels
def test_mmt4d_edge_cases_and_guards(dev, dev_ops):
    import oracle
    t, M0, N0, K0 = "bf16", 16, 16, 2
    l4 = oracle.pack(oracle.generate_random_matrix(t, (32, 8), 1), M0, K0)
    r4 = oracle.pack(oracle.generate_random_matrix(t, (16, 8), 2), N0, K0)
    out = np.full((2, 1, M0, N0), 5, dtype=np.float32)
    # K1 == 0: zeros without ACCUMULATE, untouched (and nothing launched) with it.
    dev_ops.mmt4d_host(t, 2, 1, 0, M0, N0, K0, 0, 0, out.ctypes.data)
    assert (out == 0).all()
    out[:] = 5
    launches()
    dev_ops.mmt4d_host(t, 2, 1, 0, M0, N0, K0, 0, 0, out.ctypes.data, flags=H.MATMUL_ACCUMULATE)
    dev_ops.mmt4d_host(t, 0, 1, 4, M0, N0, K0, 0, r4.ctypes.data, 0)      # empty result
    assert launches() == [] and (out == 5).all()
    for bad in (dict(M0=0), dict(K0=40000), dict(flags=H.MATMUL_TRANSPOSE_RHS), dict(t="f8e4m3fn"),
                dict(M1=-1)):
        a = dict(t=t, M1=2, N1=1, K1=4, M0=M0, N0=N0, K0=K0, flags=0)
        a.update(bad)
        with pytest.raises(H.Status):
            dev_ops.mmt4d_host(a["t"], a["M1"], a["N1"], a["K1"], a["M0"], a["N0"], a["K0"],
                               l4.ctypes.data, r4.ctypes.data, out.ctypes.data, flags=a["flags"])
    # Recorded form: operands at sub-ranges of their buffers.
    lhs_b, rhs_b = dev.allocate(l4.nbytes + 64), dev.allocate(r4.nbytes)
    dev.h2d(l4.ctypes.data, lhs_b, l4.nbytes, 64)
    dev.h2d(r4.ctypes.data, rhs_b, r4.nbytes)
    out_b = dev.allocate(out.nbytes)
    cb = H.CommandBuffer(dev).begin()
    dev_ops.record_mmt4d(cb, t, 2, 1, 4, M0, N0, K0, lhs_b.ref(64, l4.nbytes), rhs_b.ref(),
                         out_b.ref())
    cb.end()
    dev.execute(cb)
    dev.wait_idle()
    dev.d2h(out_b, out.ctypes.data, out.nbytes)
    assert np.array_equal(out, oracle.mmt4d(t, l4, r4))
    cb = H.CommandBuffer(dev).begin()
    with pytest.raises(H.Status):       # result binding too short
        dev_ops.record_mmt4d(cb, t, 2, 1, 4, M0, N0, K0, lhs_b.ref(64, l4.nbytes), rhs_b.ref(),
                             out_b.ref(0, out.nbytes - 4))
    with pytest.raises(H.Status):       # lhs at an odd byte: not a bf16 element boundary
        dev_ops.record_mmt4d(cb, t, 2, 1, 4, M0, N0, K0, lhs_b.ref(63, l4.nbytes), rhs_b.ref(),
                             out_b.ref())


@needs_kernels
def test_mxfp4_expansion_is_exact_for_every_code_and_scale(dev, dev_ops):
    """The tensor-core MX route, first half: iree_b200_matmul_mxfp4_dequant_bf16 run on the
    host. One row per scale byte (all 256), every f4 code in every block: the bf16 it writes
    equals the oracle's f32 decode exactly (NaN for scale 0xFF), and the host call's launch
    sequence is expand(lhs), expand(rhs), bf16 tcgen05 NT GEMM (logged only: no host form)."""
    import oracle
    M, N, K = 256, 128, 128
    rng = np.random.Generator(np.random.PCG64(11))
    lhs = rng.integers(0, 256, (M, K // 2), dtype=np.uint8)
    lhs[:, :8] = np.arange(8, dtype=np.uint8) * 0x22 + 0x10     # codes 0..15 in the first block
    ls = np.repeat(np.arange(256, dtype=np.uint8)[:, None], K // 32, axis=1)
    rhs = rng.integers(0, 256, (N, K // 2), dtype=np.uint8)
    rs = rng.integers(120, 135, (N, K // 32), dtype=np.uint8)
    out = np.zeros((M, N), dtype=np.float32)
    assert dev_ops.matmul_mxfp4_workspace_size(M, N, K) == M * K * 2 + N * K * 2
    assert dev_ops.matmul_mxfp4_workspace_size(127, N, K) == 0      # below one tile: SIMT
    assert dev_ops.matmul_mxfp4_workspace_size(M, 130, K) == 0      # result pitch 520 B
    launches()
    before = dev_ops.dispatch_count
    dev_ops.matmul_mxfp4_host(M, N, K, lhs.ctypes.data, ls.ctypes.data, rhs.ctypes.data,
                              rs.ctypes.data, out.ctypes.data)
    assert dev_ops.dispatch_count - before == 3
    got = launch_args()
    assert [g[0] for g in got] == ["iree_b200_matmul_mxfp4_dequant_bf16"] * 2 + \
        ["iree_b200_matmul_nt_bf16_f32_1cta"]
    (_, (_, _, a_ws, rows_a)), (_, (_, _, b_ws, rows_b)), (_, (a_gemm, b_gemm, _, m_gemm)) = got
    assert (rows_a, rows_b, m_gemm) == (M, N, M) and (a_gemm, b_gemm) == (a_ws, b_ws)
    assert b_ws == a_ws + M * K * 2 and a_ws % 256 == 0
    for ptr, packed, scales, rows in ((a_ws, lhs, ls, M), (b_ws, rhs, rs, N)):
        bits = np.ctypeslib.as_array(ctypes.cast(ptr, ctypes.POINTER(ctypes.c_uint16)), (rows, K))
        want = oracle.mx_decode(packed, scales)                  # f32 [rows, K]
        have = oracle.to_f32("bf16", bits.copy())
        assert np.array_equal(have, want, equal_nan=True)
        assert np.isnan(have[scales[:, 0] == 0xFF]).all() if (scales == 0xFF).any() else True
    # Signed zeros and the subnormal end survive: row 0 has scale 2^-127.
    row0 = np.ctypeslib.as_array(ctypes.cast(a_ws, ctypes.POINTER(ctypes.c_uint16)), (M, K))[0]
    assert set(np.unique(row0[:16])) >= {0x0000, 0x8000, 0x0020}     # +0, -0, 0.5 * 2^-127


@needs_kernels
def test_mxfp4_tensor_route_recorded_form_and_guards(dev, dev_ops):
    M, N, K = 128, 128, 128
    z = lambda n: dev.allocate(n)
    lhs, ls, rhs, rs = z(M * K // 2), z(M * K // 32), z(N * K // 2), z(N * K // 32)
    out = z(M * N * 4)
    need = dev_ops.matmul_mxfp4_workspace_size(M, N, K)
    ws = z(need + 256)
    launches()
    cb = H.CommandBuffer(dev).begin()
    dev_ops.record_matmul_mxfp4_tensor(cb, M, N, K, lhs.ref(), ls.ref(), rhs.ref(), rs.ref(),
                                       ws.ref(256, need), out.ref(), flags=H.MATMUL_ACCUMULATE)
    # No workspace: the single SIMT dispatch.
    dev_ops.record_matmul_mxfp4_tensor(cb, M, N, K, lhs.ref(), ls.ref(), rhs.ref(), rs.ref(),
                                       None, out.ref())
    with pytest.raises(H.Status):       # workspace too small
        dev_ops.record_matmul_mxfp4_tensor(cb, M, N, K, lhs.ref(), ls.ref(), rhs.ref(), rs.ref(),
                                           ws.ref(0, need - 2), out.ref())
    with pytest.raises(H.Status):       # workspace not 256-byte aligned
        dev_ops.record_matmul_mxfp4_tensor(cb, M, N, K, lhs.ref(), ls.ref(), rhs.ref(), rs.ref(),
                                           ws.ref(16, need), out.ref())
    cb.end()
    dev.execute(cb)
    dev.wait_idle()
    names = [l[0] for l in launches()]
    assert names == ["iree_b200_matmul_mxfp4_dequant_bf16"] * 2 + \
        ["iree_b200_matmul_nt_bf16_f32_1cta", "iree_b200_matmul_generic_mxfp4_f32"]


@needs_kernels
def test_mxfp4_block_scaled_route_packs_the_scales_into_atoms(dev, dev_ops):
    """The block-scaled tcgen05 route under the mock: the two scale repack dispatches run on
    host threads — their output must be the instruction's atom order, byte (r % 32) * 16 +
    (r % 128 / 32) * 4 + c of atom [r / 128][kb / 4], padding rows = scale 1.0 — and the GEMM
    dispatch reaches the driver with 8 constants, 5 bindings and 5 tensor maps (operands as
    bytes with the 128-byte swizzle, scale atoms as 128 x 2 boxes of u32, no swizzle)."""
    M, N, K = 300, 264, 512          # ragged row tiles on both sides
    g = np.random.Generator(np.random.PCG64(7))
    host = {"lhs": g.integers(0, 256, (M, K // 2), dtype=np.uint8),
            "ls": g.integers(100, 150, (M, K // 32), dtype=np.uint8),
            "rhs": g.integers(0, 256, (N, K // 2), dtype=np.uint8),
            "rs": g.integers(100, 150, (N, K // 32), dtype=np.uint8)}
    bufs = {k: dev.allocate(v.nbytes) for k, v in host.items()}
    for k, v in host.items():
        dev.h2d(v.ctypes.data, bufs[k], v.nbytes)
    out = dev.allocate(M * N * 4)
    need = dev_ops.matmul_mxfp4_scale_workspace_size(M, N, K)
    tiles = lambda rows: -(-rows // 128)
    assert need == (tiles(M) + tiles(N)) * (K // 128) * 512
    assert dev_ops.matmul_mxfp4_scale_workspace_size(M, N, 384) == 0     # K % 256
    assert dev_ops.matmul_mxfp4_scale_workspace_size(M, 266, K) == 0     # N % 4
    ws = dev.allocate(need + 256)
    launches(), tensor_maps()
    cb = H.CommandBuffer(dev).begin()
    dev_ops.record_matmul_mxfp4_tensor(cb, M, N, K, bufs["lhs"].ref(), bufs["ls"].ref(),
                                       bufs["rhs"].ref(), bufs["rs"].ref(), ws.ref(256, need), out.ref())
    # Without a workspace: ONE dispatch of the form that gathers the scales itself.
    dev_ops.record_matmul_mxfp4(cb, M, N, K, bufs["lhs"].ref(), bufs["ls"].ref(), bufs["rhs"].ref(),
                                bufs["rs"].ref(), out.ref())
    cb.end()
    dev.execute(cb)
    dev.wait_idle()
    names = [l[0] for l in launches()]
    # (a 3 x 2-tile problem is far below one tile pair per SM: the 1-CTA family)
    assert names == ["iree_b200_mxfp4_pack_scales"] * 2 + ["iree_b200_matmul_mxfp4_packed_f32_1cta",
                                                            "iree_b200_matmul_mxfp4_f32_1cta"]
    maps = tensor_maps()
    assert len(maps) == 5 + 3
    packed = maps[:5]
    assert [m["type"] for m in packed] == [0, 0, 7, 2, 2]            # u8, u8, f32, u32, u32
    assert packed[0]["dims"][:2] == [K // 2, M] and packed[1]["dims"][:2] == [K // 2, N]
    assert packed[3]["dims"][:3] == [128, K // 128, tiles(M)] and packed[3]["box"][:3] == [128, 2, 1]
    assert packed[4]["dims"][:3] == [128, K // 128, tiles(N)] and packed[3]["swizzle"] == 0
    atoms = np.zeros(need, dtype=np.uint8)
    dev.d2h(ws, atoms.ctypes.data, need, offset=256)
    dev.wait_idle()

    def want_atoms(scales, rows):
        a = np.full((tiles(rows), K // 128, 32, 4, 4), 0x7F, dtype=np.uint8)   # [rt][ka][r%32][r%128/32][c]
        for r in range(rows):
            a[r // 128, :, r % 32, (r % 128) // 32, :] = scales[r].reshape(K // 128, 4)
        return a.reshape(-1)

    split = tiles(M) * (K // 128) * 512
    assert np.array_equal(atoms[:split], want_atoms(host["ls"], M))
    assert np.array_equal(atoms[split:], want_atoms(host["rs"], N))


@needs_kernels
@pytest.mark.parametrize("t,tiles", [("bf16", (16, 16, 2)), ("i8", (8, 32, 4)), ("f16", (16, 16, 16))],
                         ids=str)
def test_mmt4d_tensor_route_relayouts(dev, dev_ops, t, tiles):
    """The tensor-core mmt4d route under the mock: the relayout kernels run on the host, the
    tcgen05 GEMM in the middle is only logged. Checked: the dispatch sequence, that the GEMM's
    operand bindings hold exactly the unpacked (row-major) lhs / rhs, and — with ACCUMULATE —
    that accumulator -> row-major -> tiles is the identity (the GEMM being a no-op here)."""
    import oracle
    M0, N0, K0 = tiles
    M, K, N = 128, 128, 256
    lhs = oracle.generate_random_matrix(t, (M, K), 1)
    rhs = oracle.generate_random_matrix(t, (K, N), 2)
    l4 = oracle.pack(lhs, M0, K0)
    r4 = oracle.pack(np.ascontiguousarray(rhs.T), N0, K0)
    M1, K1, N1 = l4.shape[0], l4.shape[1], r4.shape[0]
    acc_t = "i32" if t == "i8" else "f32"
    acc = oracle.generate_random_matrix(acc_t, (M1, N1, M0, N0), 3)
    es = l4.dtype.itemsize
    need = dev_ops.mmt4d_workspace_size(t, M1, N1, K1, M0, N0, K0)
    assert need == M * K * es + N * K * es + M * N * 4
    assert dev_ops.mmt4d_workspace_size(t, M1 - 1, N1, K1, M0, N0, K0) == 0    # < 128 rows
    out = acc.copy()
    launches()
    before = dev_ops.dispatch_count
    dev_ops.mmt4d_host(t, M1, N1, K1, M0, N0, K0, l4.ctypes.data, r4.ctypes.data, out.ctypes.data,
                       flags=H.MATMUL_ACCUMULATE)
    assert dev_ops.dispatch_count - before == 5
    got = launch_args()
    relayout = {1: "iree_b200_mmt4d_relayout_b8", 2: "iree_b200_mmt4d_relayout_b16"}[es]
    names = [g[0] for g in got]
    assert names[:3] == [relayout, relayout, "iree_b200_mmt4d_relayout_b32"]
    assert names[3].startswith("iree_b200_matmul_nt_%s_%s" % (t, acc_t))
    assert names[4] == "iree_b200_mmt4d_relayout_b32"
    a_rm, b_rm = got[3][1][0], got[3][1][1]
    ct = {1: ctypes.c_int8, 2: ctypes.c_uint16}[es]
    a_have = np.ctypeslib.as_array(ctypes.cast(a_rm, ctypes.POINTER(ct)), (M, K))
    b_have = np.ctypeslib.as_array(ctypes.cast(b_rm, ctypes.POINTER(ct)), (N, K))
    assert np.array_equal(a_have.view(lhs.dtype), lhs)
    assert np.array_equal(b_have.view(rhs.dtype), np.ascontiguousarray(rhs.T))
    assert np.array_equal(out, acc)
    # Without ACCUMULATE the accumulator is not unpacked: 4 dispatches.
    before = dev_ops.dispatch_count
    dev_ops.mmt4d_host(t, M1, N1, K1, M0, N0, K0, l4.ctypes.data, r4.ctypes.data, out.ctypes.data)
    assert dev_ops.dispatch_count - before == 4
    launches()
