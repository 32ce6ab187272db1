"""Multi-rank GPU parity inside `pytest -m gpu` (skipped on a 1-GPU box): one process per GPU
under torchrun, as the product is launched. Mirrors the per-rank-count structure of the
reference's tests/e2e/collectives/collectives_test.py:46-93.

* tools/symmetric_check.py — all_reduce(sum) through iree_b200_command_buffer_collective on
  our NVLink kernels (NVLS + peer load/store), f16/bf16/f32, ragged sizes, sub-ranges,
  back-to-back launches, fused ranges, NCCL fallback: exact sums.
* tools/collectives_check.py — every collective kind through NCCL on real ranks, stream and
  graph-captured, exact.
* tools/tp_layer_check.py — the tensor-parallel decoder-layer GEMMs against the oracle's
  unsharded layer, bit for bit, through NCCL (stream / overlapped / graph-captured) and
  through the symmetric kernels (staged / fused / graph kernel nodes).
"""

import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def device_count():
    import torch
    return torch.cuda.device_count()


def torchrun(world, script, *script_args, port):
    env = dict(os.environ, IREE_B200_COLLECTIVE_WATCHDOG_MS="30000")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, script),
           *script_args]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, env=env, cwd=ROOT)
    lines = [json.loads(l) for l in r.stdout.splitlines() if l.startswith("{")]
    return r, lines


@pytest.mark.parametrize("world", [2, 4, 8])
def test_symmetric_all_reduce_ranks(world):
    if device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    r, lines = torchrun(world, "tools/symmetric_check.py", "--mib", "16", "--iters", "3",
                        port=29560 + world)
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-1500:])
    assert lines and lines[-1]["world"] == world and lines[-1]["failures"] == []
    assert lines[-1]["failing_ranks_total"] == 0


@pytest.mark.parametrize("world", [2, 4, 8])
def test_every_collective_kind_on_real_ranks(world):
    """tools/collectives_check.py: all_reduce (sum / max), all_gather, reduce_scatter, broadcast,
    reduce, all_to_all, send_recv, send / recv through iree_b200_command_buffer_collective, f32 /
    bf16 / i32, in stream mode and captured into a CUgraph (collectives_test.py:46-93)."""
    if device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    r, lines = torchrun(world, "tools/collectives_check.py", port=29580 + world)
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-1500:])
    assert lines and lines[-1]["world"] == world and lines[-1]["failures"] == []
    assert lines[-1]["cases"] >= 2 * (3 * 7 + 1)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_tensor_parallel_decoder_layer_matches_the_oracle(world):
    if device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    r, lines = torchrun(world, "tools/tp_layer_check.py", port=29570 + world)
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-1500:])
    assert lines and lines[-1]["world"] == world and lines[-1]["failures"] == []
    assert len(lines[-1]["paths"]) == 16
