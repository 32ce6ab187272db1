"""An IREE MODULE around our kernels, run by the reference's own stock tools (SURVEY.md §7.1 step
7, the `.vmasm` route): iree_b200/vmasm.py writes the host half the compiler would emit — HAL
calls in VM assembly with a CDA1 executable as rodata — the reference's `iree-as-module` turns
it into a .vmfb, and the UNMODIFIED `iree-run-module --device=cuda` loads and runs it: stock
VM, stock HAL module, stock CUDA driver, our tcgen05 kernels (the `*_dtm` exports, which build
their tensor maps on the device).

* CPU: the assembly assembles, the module passes the reference's bytecode verifier, declares
  its `hal` dependency, and on the CPU driver its initialiser runs up to hal.executable.create,
  which rejects the CUDA executable format — everything before the GPU is exercised.
* GPU: results of iree-run-module compared with the oracle, bit for bit for the matmuls.
"""

import os
import struct
import subprocess

import numpy as np
import pytest

import oracle
from iree_b200 import vmasm

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
STOCK = os.path.join(ROOT, "oracle", "_ref", "iree-rt", "stock")


def tool(name):
    path = os.path.join(STOCK, name)
    if not os.path.exists(path):
        pytest.skip(f"{path} not built (tools/build_ref_runtime.sh needs /root/reference)")
    return path


@pytest.fixture(scope="module")
def module_path():
    if os.path.exists(vmasm.MODULE_VMFB) and not os.path.exists(vmasm.DEFAULT_ASSEMBLER):
        return vmasm.MODULE_VMFB            # GPU box: built on the CPU box, shipped
    tool("iree-as-module")
    return vmasm.build_module()


def test_module_assembles_verifies_and_declares_its_dependency(module_path):
    text = subprocess.run([tool("iree-dump-module"), module_path], capture_output=True, text=True,
                          check=True).stdout
    assert "hal, version >= 0, required" in text
    for w in vmasm.default_workloads():
        assert f"] {w.name}(" in text and f"] {w.name}_into(" in text
    assert "`CDA1`" in text and "`PTXE`" in text
    for name, _ in vmasm.IMPORTS:
        assert f"] {name}(" in text


def test_dependency_splice_keeps_the_module_intact(tmp_path):
    """add_dependencies on the assembler's raw output: same size-prefixed container, every root
    field still resolves to the same bytes, one new vector."""
    asm = tmp_path / "m.vmasm"
    w = vmasm.default_workloads()[4:5]
    asm.write_text(vmasm.generate(w))
    raw_path = tmp_path / "raw.vmfb"
    subprocess.run([tool("iree-as-module"), f"--output={raw_path}", str(asm)], check=True)
    raw = raw_path.read_bytes()
    out = vmasm.add_dependencies(raw, ("hal",))
    assert struct.unpack_from("<I", out, 0)[0] == len(out) - 4 and out[8:12] == b"IREE"
    assert (len(out) - len(raw)) % 64 == 0 and out[-len(raw) + 12:] == raw[12:]
    with pytest.raises(ValueError):
        vmasm.add_dependencies(out)                      # already declared
    with pytest.raises(ValueError):
        vmasm.add_dependencies(raw[:-4])                 # truncated
    fixed = tmp_path / "fixed.vmfb"
    fixed.write_bytes(out)
    for path, want in ((raw_path, False), (fixed, True)):
        text = subprocess.run([tool("iree-dump-module"), str(path)], capture_output=True,
                              text=True, check=True).stdout
        assert ("hal, version >= 0" in text) == want
        assert "matmul_bf16_small" in text


def test_stock_tool_runs_the_initialiser_up_to_the_cuda_executable(module_path):
    r = subprocess.run([tool("iree-run-module"), "--device=local-sync", f"--module={module_path}",
                        "--function=matmul_bf16_small", "--input=1000x264xbf16=1",
                        "--input=264x520xbf16=1"], capture_output=True, text=True)
    assert r.returncode != 0
    assert "no executable loader registered for the given executable format 'PTXE'" in r.stderr
    assert "hal.executable.create" in r.stderr


def _run(module_path, tmp_path, function, inputs, out_dtype, out_shape):
    args = [tool("iree-run-module"), "--device=cuda", f"--module={module_path}",
            f"--function={function}"]
    for i, (a, t) in enumerate(inputs):
        f = tmp_path / f"in{i}.bin"
        a.tofile(f)
        args.append("--input=%sx%s=@%s" % ("x".join(str(d) for d in a.shape), t, f))
    out = tmp_path / "out.bin"
    args.append(f"--output=@{out}")
    r = subprocess.run(args, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    return np.fromfile(out, dtype=out_dtype).reshape(out_shape)


@pytest.mark.gpu
@pytest.mark.parametrize("t,function,shape", [("bf16", "matmul_bf16_small", (1000, 264, 520)),
                                              ("i8", "matmul_i8_small", (1000, 272, 528))])
def test_stock_iree_run_module_runs_the_tcgen05_matmul(module_path, tmp_path, t, function, shape):
    M, K, N = shape
    lhs = oracle.generate_random_matrix(t, (M, K), 1)
    rhs = oracle.generate_random_matrix(t, (K, N), 2)
    acc_t, acc = ("i32", np.int32) if t == "i8" else ("f32", np.float32)
    got = _run(module_path, tmp_path, function, [(lhs, t), (rhs, t)], acc, (M, N))
    want = oracle.matmul(t, lhs, rhs)
    assert np.array_equal(got, want)
    # Destination passing: the caller's buffer view comes back holding the result.
    into = _run(module_path, tmp_path, function + "_into",
                [(lhs, t), (rhs, t), (np.full((M, N), 77, dtype=acc), acc_t)], acc, (M, N))
    assert np.array_equal(into, want)


@pytest.mark.gpu
def test_stock_iree_run_module_runs_the_fused_linear_layer(module_path, tmp_path):
    """matmul + bias[n] + relu + truncf as one tcgen05 dispatch (`iree_b200_matmul_bias_bf16_bf16_
    1cta_dtm`), four bindings, through the stock tool. Bit-exact."""
    M, K, N = 1000, 264, 520
    lhs = oracle.generate_random_matrix("bf16", (M, K), 1)
    rhs = oracle.generate_random_matrix("bf16", (K, N), 2)
    bias = oracle.generate_random_matrix("f32", (N,), 4)
    got = _run(module_path, tmp_path, "linear_bf16_small", [(lhs, "bf16"), (rhs, "bf16"), (bias, "f32")],
               np.uint16, (M, N))
    assert np.array_equal(got, oracle.matmul_bias("bf16", lhs, rhs, bias, relu=True, truncate=True))


@pytest.mark.gpu
def test_stock_iree_run_module_runs_attention(module_path, tmp_path):
    g = np.random.Generator(np.random.PCG64(0xB200))
    mk = lambda *s: oracle.from_f32("f16", np.clip(g.standard_normal(s), -4, 4).astype(np.float32))
    scale = 1.0 / np.sqrt(128.0)
    # tcgen05 prefill with the causal mask computed in the kernel, tensor maps built on the device.
    B, M, K2 = 2, 384, 640
    q, k, v = mk(B, M, 128), mk(B, K2, 128), mk(B, K2, 128)
    got = _run(module_path, tmp_path, "attention_prefill_causal_f16_small",
               [(q, "f16"), (k, "f16"), (v, "f16")], np.uint16, (B, M, 128))
    mask = np.broadcast_to(np.arange(K2)[None, :] <= np.arange(M)[:, None] + (K2 - M), (B, M, K2)).copy()
    want = oracle.to_f32("f16", oracle.attention("f16", q, k, v, scale, mask=mask, k2_tile=128))
    assert np.abs(oracle.to_f32("f16", got) - want).max() <= 2e-3 * np.abs(want).max()
    # The Llama decode shape on the 8-CTA cluster kernel.
    B, M, K2 = 32, 1, 8192
    q, k, v = mk(B, M, 128), mk(B, K2, 128), mk(B, K2, 128)
    got = _run(module_path, tmp_path, "attention_decode_f16", [(q, "f16"), (k, "f16"), (v, "f16")],
               np.uint16, (B, M, 128))
    want = oracle.to_f32("f16", oracle.attention("f16", q, k, v, scale, k2_tile=128))
    assert np.abs(oracle.to_f32("f16", got) - want).max() <= 2e-3 * np.abs(want).max()
