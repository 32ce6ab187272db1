#!/usr/bin/env python3
"""Extracts the attention known-answer vectors from the reference's check tests into
tests/golden/attention_golden.json.

Sources (read-only, only available in the build container):
  /root/reference/tests/e2e/linalg_ext_ops/attention.mlir
  /root/reference/tests/e2e/linalg_ext_ops/attention_i1_mask.mlir
  /root/reference/tests/e2e/linalg_ext_ops/dynamic_attention.mlir

Each `func.func` holds dense constants for query/key/value (+ optional i1 mask, given
either as a bool tensor or as packed bytes bit-cast to i1), a scalar scale, an optional
softcap region, and the expected output of `check.expect_almost_eq_const`
(atol=1e-4, compiler/src/iree/compiler/Modules/Check/IR/CheckOps.td:163-164).

Run once here; the JSON is committed because /root/reference does not exist on the
GPU box.
"""

import json
import os
import re
import sys

REF = "/root/reference/tests/e2e/linalg_ext_ops"
FILES = ["attention.mlir", "attention_i1_mask.mlir", "dynamic_attention.mlir"]
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden",
                   "attention_golden.json")


def parse_dense(text, pos):
    """Parses `dense<...>` starting at text[pos:] == 'dense<'. Returns (py, end)."""
    assert text.startswith("dense<", pos)
    i = pos + len("dense<")
    depth = 0
    j = i
    while True:
        c = text[j]
        if c == "[":
            depth += 1
        elif c == "]":
            depth -= 1
        elif c == ">" and depth == 0:
            break
        j += 1
    body = text[i:j]
    body = body.replace("true", "1").replace("false", "0")
    return json.loads(body), j + 1


def tensor_shape_after(text, pos):
    m = re.compile(r"\s*:\s*tensor<([0-9x]+)x(f32|i1|i8)").match(text, pos)
    assert m, text[pos:pos + 80]
    return [int(d) for d in m.group(1).split("x")], m.group(2)


def extract_func(name, body, src):
    entry = {"name": name, "source": src}
    # Named constants.
    for m in re.finditer(r"%(\w+)\s*=\s*(?:util\.unfoldable_constant|flow\.tensor\.dynamic_constant)\s+", body):
        var = m.group(1)
        p = m.end()
        if not body.startswith("dense<", p):
            continue
        val, end = parse_dense(body, p)
        shape, ety = tensor_shape_after(body, end)
        entry[var] = {"shape": shape, "type": ety, "data": val}
    m = re.search(r"%scale\s*=\s*arith\.constant\s+([0-9.eE+-]+)\s*:\s*f32", body)
    entry["scale"] = float(m.group(1))
    # Packed i1 mask: `flow.tensor.bitcast %i8mask : tensor<2xi8> -> tensor<1x4x4xi1, ...packed_storage>`.
    m = re.search(r"%mask\s*=\s*flow\.tensor\.bitcast\s+%i8mask.*->\s*tensor<([0-9x]+)xi1", body)
    if m:
        shape = [int(d) for d in m.group(1).split("x")]
        raw = entry.pop("i8mask")["data"]
        n = 1
        for d in shape:
            n *= d
        # packed_storage: element i lives in bit (i % 8) of byte i // 8 (LSB first).
        bits = [((raw[i // 8] & 0xFF) >> (i % 8)) & 1 for i in range(n)]
        entry["mask"] = {"shape": shape, "type": "i1", "data": bits, "packed_bytes": raw}
    # Softcap region: tanh(x) * cst.
    if "math.tanh" in body:
        m = re.search(r"%cst\s*=\s*arith\.constant\s+([0-9.eE+-]+)\s*:\s*f32\s*\n\s*%mul\s*=\s*arith\.mulf\s+%tanh", body)
        entry["region"] = {"kind": "softcap", "param": float(m.group(1))}
    else:
        entry["region"] = {"kind": "identity", "param": 0.0}
    entry["op"] = "online_attention" if "iree_linalg_ext.online_attention" in body else "attention"
    # Expected.
    m = re.search(r"check\.expect_almost_eq_const\(\s*%\w+,\s*", body)
    val, end = parse_dense(body, m.end())
    shape, ety = tensor_shape_after(body, end)
    entry["expected"] = {"shape": shape, "type": ety, "data": val}
    entry["atol"] = 1e-4
    return entry


def main():
    cases = []
    for f in FILES:
        text = open(os.path.join(REF, f)).read()
        parts = re.split(r"(?m)^func\.func @(\w+)\(\)\s*\{", text)
        # parts = [pre, name1, body1, name2, body2, ...]
        for i in range(1, len(parts), 2):
            cases.append(extract_func(parts[i], parts[i + 1], f))
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    with open(OUT, "w") as fh:
        json.dump({"generator": "tools/extract_attention_golden.py",
                   "reference_commit": "b48a5955575cf8924ffc191f3b0804976e5738b9",
                   "cases": cases}, fh, indent=1)
    print(f"wrote {len(cases)} cases to {os.path.normpath(OUT)}")
    for c in cases:
        print("  ", c["source"], c["name"], c["op"], "mask" if "mask" in c else "",
              c["region"]["kind"])


if __name__ == "__main__":
    sys.exit(main())
