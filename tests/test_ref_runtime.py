"""The drop-in boundary run through the REFERENCE's own runtime.

tools/build_ref_runtime.sh builds, from /root/reference on the CPU box, the reference runtime
(HAL core + its CUDA HAL driver) with two checkers of ours linked against it, into
oracle/_ref/iree-rt/ (git-ignored, shipped to the GPU box):

* CPU, no GPU: `cda1_ref_verify` feeds the CDA1 containers `iree_b200/cda1.py` writes to the
  reference's header parser and flatcc-generated verifier — the exact front end of
  `iree_hal_cuda_native_executable_create` (native_executable.c:141-150) — and reads every
  export back with the reference's generated reader.
* GPU: `ref_hal_harness` loads the container through `iree_hal_executable_cache_prepare_
  executable` of the reference driver, dispatches one export through
  `iree_hal_command_buffer_dispatch` / `iree_hal_device_queue_execute`
  (stream_command_buffer.c:481-609) and writes the result binding to a file; the result is
  compared with the oracle, bit for bit. `stock/` is the unmodified driver (recipe-free
  exports); `tma/` is the driver with integration/cuda_driver_tma.patch (tcgen05 exports).
"""

import json
import os
import subprocess

import numpy as np
import pytest

import oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RT = os.path.join(ROOT, "oracle", "_ref", "iree-rt")
BUILD = os.path.join(ROOT, "iree_b200", "_build")
CDA1 = os.path.join(BUILD, "iree_b200_kernels.cda1")


def tool(flavour, name):
    path = os.path.join(RT, flavour, name)
    if not os.path.exists(path):
        pytest.skip(f"{path} not built (tools/build_ref_runtime.sh needs /root/reference)")
    return path


def verify(path):
    r = subprocess.run([tool("stock", "cda1_ref_verify"), path], capture_output=True, text=True)
    return r.returncode, [json.loads(l) for l in r.stdout.splitlines() if l.startswith("{")], r.stderr


@pytest.mark.parametrize("name", ["iree_b200_kernels.cda1", "iree_b200_pending_kernels.cda1"])
def test_reference_loader_front_end_accepts_our_cda1(name):
    path = os.path.join(BUILD, name)
    if not os.path.exists(path):
        pytest.skip("executable not built")
    code, lines, err = verify(path)
    assert code == 0, err
    head = next(l for l in lines if "exports" in l)
    assert head["file_size"] == os.path.getsize(path)
    exports = [l for l in lines if "ordinal" in l]
    assert head["exports"] == len(exports) > 0
    with open(os.path.join(BUILD, "exports.json")) as f:
        manifest = {e["name"]: e for e in json.load(f)["exports"]}
    # The reference's reader sees exactly what the writer's manifest says.
    for got in exports:
        want = manifest[got["kernel_name"]]
        assert got["block_dims"] == want["block_dims"]
        assert got["block_shared_memory_size"] == want["shared_memory"]
        assert got["constant_count"] == want["constants"]
        assert got["binding_count"] == len(want["binding_flags"])
        # Loader limits (native_executable.h:24,28; :192-221 with B200's 227 KiB opt-in).
        assert got["constant_count"] <= 64 and got["binding_count"] <= 16
        assert got["block_shared_memory_size"] <= 232448
    for m in (l for l in lines if "image_bytes" in l):
        assert m["image_bytes"] > 0


def test_reference_loader_front_end_rejects_damaged_containers(tmp_path):
    if not os.path.exists(CDA1):
        pytest.skip("executable not built")
    blob = bytearray(open(CDA1, "rb").read())
    cases = {
        "magic": bytes([blob[0] ^ 0xFF]) + bytes(blob[1:]),
        "truncated": bytes(blob[: len(blob) // 2]),
        "file_id": bytes(blob[:68]) + b"XXXX" + bytes(blob[72:]),
        "root_offset": bytes(blob[:64]) + (0x7FFFFFF0).to_bytes(4, "little") + bytes(blob[68:]),
    }
    for label, data in cases.items():
        p = tmp_path / f"{label}.cda1"
        p.write_bytes(data)
        code, _, err = verify(str(p))
        assert code != 0, label
        # ... and our own reader agrees with the reference on every one of them.
    from iree_b200 import hal as H
    for label, data in cases.items():
        with pytest.raises(H.Status):
            H.infer_format(data)


# ---------------------------------------------------------------------------------------
# GPU: dispatch through the reference HAL
# ---------------------------------------------------------------------------------------

def run_harness(flavour, tmp_path, ordinal, grid, constants, bindings, out_index, iterations=0,
                batch_size=1):
    """bindings: list of np.ndarray (inputs, written to files) or int (zero-filled bytes)."""
    cmd = [tool(flavour, "ref_hal_harness"), "--device=cuda", f"--executable_file={CDA1}",
           f"--entry_point={ordinal}", "--workgroup_count=%d,%d,%d" % tuple(grid),
           f"--batch_size={batch_size}", f"--iterations={iterations}"]
    cmd += [f"--constant={int(c)}" for c in constants]
    nbytes = []
    for i, b in enumerate(bindings):
        if isinstance(b, np.ndarray):
            f = tmp_path / f"binding{i}.bin"
            b.tofile(f)
            cmd.append(f"--binding={b.nbytes}=@{f}")
            nbytes.append(b.nbytes)
        else:
            cmd.append(f"--binding={int(b)}")
            nbytes.append(int(b))
    out_file = tmp_path / "out.bin"
    cmd.append(f"--output={out_index}={out_file}")
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [json.loads(l) for l in r.stdout.splitlines() if l.startswith("{")]
    assert lines[0]["loaded"] is True
    return np.fromfile(out_file, dtype=np.uint8), lines


@pytest.mark.gpu
@pytest.mark.parametrize("t,shape", [("bf16", (457, 330, 512)), ("i8", (513, 128, 55)),
                                     ("f16", (130, 72, 264))])
def test_stock_reference_driver_runs_the_generic_matmul(ops, hal, tmp_path, t, shape):
    M, K, N = shape
    sel = ops.select_matmul(t, 1, M, N, K, flags=hal.MATMUL_FORCE_GENERIC)
    name = ops.executable.export_info(sel["export_ordinal"])["kernel_name"]
    assert name.startswith("iree_b200_matmul_generic_")
    lhs = oracle.generate_random_matrix(t, (M, K), 1)
    rhs = oracle.generate_random_matrix(t, (K, N), 2)
    raw, _ = run_harness("stock", tmp_path, sel["export_ordinal"], sel["workgroup_count"],
                         sel["constants"], [lhs, rhs, M * N * 4], 2)
    out = raw.view(np.int32 if t == "i8" else np.float32).reshape(M, N)
    assert np.array_equal(out, oracle.matmul(t, lhs, rhs))


@pytest.mark.gpu
@pytest.mark.parametrize("t,shape,flags", [
    ("bf16", (512, 512, 512), 0), ("i8", (1000, 272, 528), 0), ("f16", (384, 768, 1024), 0),
    ("bf16", (2048, 1280, 2048), 0), ("i8", (4096, 512, 2048), "nt")])
def test_stock_reference_driver_runs_the_tcgen05_matmul_with_device_built_maps(ops, hal, tmp_path, t,
                                                                               shape, flags):
    """The `*_dtm` exports: tcgen05 + TMA kernels that build their tensor maps on the device
    (tensormap.replace on host-encoded templates), so the UNMODIFIED reference driver — binding
    pointers and push constants only — launches them. Bit-exact."""
    M, K, N = shape
    nt = flags == "nt"
    sel = ops.select_matmul(t, 1, M, N, K, flags=hal.MATMUL_DEVICE_MAPS |
                            (hal.MATMUL_TRANSPOSE_RHS if nt else 0))
    info = ops.executable.export_info(sel["export_ordinal"])
    assert "_dtm" in info["kernel_name"] and info["tma_map_count"] == 0
    lhs = oracle.generate_random_matrix(t, (M, K), 1)
    rhs = oracle.generate_random_matrix(t, (K, N), 2)
    raw, _ = run_harness("stock", tmp_path, sel["export_ordinal"], sel["workgroup_count"],
                         sel["constants"], [lhs, np.ascontiguousarray(rhs.T) if nt else rhs, M * N * 4], 2)
    out = raw.view(np.int32 if t == "i8" else np.float32).reshape(M, N)
    assert np.array_equal(out, oracle.matmul(t, lhs, rhs))


@pytest.mark.gpu
@pytest.mark.parametrize("t", ["bf16", "f8e4m3fn"])
def test_stock_reference_driver_runs_the_other_device_map_twins(ops, hal, tmp_path, t):
    """f8 operands, and the fused bias + relu + truncf dispatch (four bindings), as `*_dtm`
    exports through the unmodified reference driver."""
    M, K, N = 1000, 272, 528
    lhs = oracle.generate_random_matrix(t, (M, K), 1)
    rhs = oracle.generate_random_matrix(t, (K, N), 2)
    if t == "bf16":
        flags = hal.MATMUL_BIAS | hal.MATMUL_RELU | hal.MATMUL_TRUNCATE_RESULT | hal.MATMUL_DEVICE_MAPS
        bias = oracle.generate_random_matrix("f32", (N,), 4)
        bindings, out_index, out_bytes = [lhs, rhs, bias, M * N * 2], 3, 2
        want = oracle.matmul_bias(t, lhs, rhs, bias, relu=True, truncate=True)
    else:
        flags = hal.MATMUL_DEVICE_MAPS
        bindings, out_index, out_bytes = [lhs, rhs, M * N * 4], 2, 4
        want = oracle.matmul(t, lhs, rhs)
    sel = ops.select_matmul(t, 1, M, N, K, flags=flags)
    info = ops.executable.export_info(sel["export_ordinal"])
    assert info["kernel_name"].endswith("_dtm") and info["tma_map_count"] == 0
    raw, _ = run_harness("stock", tmp_path, sel["export_ordinal"], sel["workgroup_count"],
                         sel["constants"], bindings, out_index)
    got = raw.view(np.uint16 if out_bytes == 2 else np.float32).reshape(M, N)
    assert np.array_equal(got, want)


@pytest.mark.gpu
def test_stock_reference_driver_runs_the_decode_cluster_attention(ops, hal, tmp_path):
    """One-dispatch split-KV decode: 8-CTA clusters + 1-D bulk copies, no tensor map, so the
    unmodified reference driver's cuLaunchKernel is enough."""
    B, M, K2, D = 32, 1, 8192, 128
    g = np.random.Generator(np.random.PCG64(0xB200))
    mk = lambda *s: oracle.from_f32("f16", np.clip(g.standard_normal(s), -4, 4).astype(np.float32))
    q, k, v = mk(B, M, D), mk(B, K2, D), mk(B, K2, D)
    scale = 1.0 / np.sqrt(D)
    sel = ops.select_attention("f16", B, M, D, K2, D, scale)
    assert len(sel["dispatches"]) == 1
    d = sel["dispatches"][0]
    name = ops.executable.export_info(d["export_ordinal"])["kernel_name"]
    assert name == "iree_b200_attention_decode_cluster_f16"
    raw, _ = run_harness("stock", tmp_path, d["export_ordinal"], d["workgroup_count"],
                         d["constants"], [q, k, v, B * M * D * 2], 3)
    got = oracle.to_f32("f16", raw.view(np.uint16).reshape(B, M, D))
    want = oracle.to_f32("f16", oracle.attention("f16", q, k, v, scale, k2_tile=128))
    assert np.abs(got - want).max() <= 2e-3 * np.abs(want).max()


@pytest.mark.gpu
@pytest.mark.parametrize("t,shape", [("bf16", (512, 512, 512)), ("i8", (1000, 272, 528)),
                                     ("f16", (384, 768, 1024)), ("bf16", (2048, 1280, 256))])
def test_patched_reference_driver_runs_the_tcgen05_matmul(ops, hal, tmp_path, t, shape):
    """The TMA-fed tcgen05 exports through the reference driver + the 60-line hook of
    integration/cuda_driver_tma.patch: recipe read at load, tensor maps encoded at dispatch."""
    M, K, N = shape
    sel = ops.select_matmul(t, 1, M, N, K)
    name = ops.executable.export_info(sel["export_ordinal"])["kernel_name"]
    assert "generic" not in name
    lhs = oracle.generate_random_matrix(t, (M, K), 1)
    rhs = oracle.generate_random_matrix(t, (K, N), 2)
    raw, _ = run_harness("tma", tmp_path, sel["export_ordinal"], sel["workgroup_count"],
                         sel["constants"], [lhs, rhs, M * N * 4], 2)
    out = raw.view(np.int32 if t == "i8" else np.float32).reshape(M, N)
    assert np.array_equal(out, oracle.matmul(t, lhs, rhs))


@pytest.mark.gpu
def test_patched_reference_driver_runs_the_tcgen05_prefill_attention(ops, hal, tmp_path):
    B, M, K2, D = 2, 512, 640, 128
    g = np.random.Generator(np.random.PCG64(0xB200))
    mk = lambda *s: oracle.from_f32("f16", np.clip(g.standard_normal(s), -4, 4).astype(np.float32))
    q, k, v = mk(B, M, D), mk(B, K2, D), mk(B, K2, D)
    scale = 1.0 / np.sqrt(D)
    sel = ops.select_attention("f16", B, M, D, K2, D, scale, flags=hal.ATTENTION_NO_TAIL_SPLIT)
    assert len(sel["dispatches"]) == 1
    d = sel["dispatches"][0]
    assert ops.executable.export_info(d["export_ordinal"])["kernel_name"] == "iree_b200_attention_prefill_f16"
    raw, _ = run_harness("tma", tmp_path, d["export_ordinal"], d["workgroup_count"],
                         d["constants"], [q, k, v, B * M * D * 2], 3)
    got = oracle.to_f32("f16", raw.view(np.uint16).reshape(B, M, D))
    want = oracle.to_f32("f16", oracle.attention("f16", q, k, v, scale, k2_tile=128))
    assert np.abs(got - want).max() <= 2e-3 * np.abs(want).max()


@pytest.mark.gpu
@pytest.mark.parametrize("t,causal", [("f16", False), ("bf16", True)])
def test_stock_reference_driver_runs_the_tcgen05_prefill_attention(ops, hal, tmp_path, t, causal):
    """`iree_b200_attention_prefill_*_dtm`: the tcgen05 prefill kernel building its Q / K / V
    tensor maps on the device, launched by the UNMODIFIED reference driver."""
    B, M, K2, D = 3, 640, 640, 128
    g = np.random.Generator(np.random.PCG64(0xB200))
    mk = lambda *s: oracle.from_f32(t, np.clip(g.standard_normal(s), -4, 4).astype(np.float32))
    q, k, v = mk(B, M, D), mk(B, K2, D), mk(B, K2, D)
    scale = 1.0 / np.sqrt(D)
    sel = ops.select_attention(t, B, M, D, K2, D, scale, flags=hal.ATTENTION_DEVICE_MAPS,
                               mask_kind=hal.ATTENTION_MASK_CAUSAL if causal else hal.ATTENTION_MASK_NONE)
    assert len(sel["dispatches"]) == 1
    d = sel["dispatches"][0]
    info = ops.executable.export_info(d["export_ordinal"])
    assert info["kernel_name"] == f"iree_b200_attention_prefill_{t}_dtm" and info["tma_map_count"] == 0
    raw, _ = run_harness("stock", tmp_path, d["export_ordinal"], d["workgroup_count"],
                         d["constants"], [q, k, v, B * M * D * 2], 3)
    got = oracle.to_f32(t, raw.view(np.uint16).reshape(B, M, D))
    mask = np.tril(np.ones((M, K2), dtype=bool))[None].repeat(B, 0) if causal else None
    want = oracle.to_f32(t, oracle.attention(t, q, k, v, scale, k2_tile=128, mask=mask))
    assert np.abs(got - want).max() <= (2e-2 if t == "bf16" else 2e-3) * np.abs(want).max()
