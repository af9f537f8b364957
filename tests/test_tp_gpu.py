"""GPU-side tensor-parallel checks, launched under torchrun from pytest (skipped below 2 GPUs; the driver's
SCALE-sized leases run them at 2, 4 and 8 ranks): tests/multi_gpu_tp.py (the unchanged scheduler + model sharded by
`tp_runtime`, NCCL and K8 all-reduce, TP MLP, vs the unsharded drop-in stack) and tests/multi_gpu_oproj.py (K8 alone,
incl. CUDA-graph replay)."""
import os
import subprocess
import sys

import pytest
import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worlds():
    n = torch.cuda.device_count() if torch.cuda.is_available() else 0
    return [w for w in (2, 4, 8) if w <= n]


def _torchrun(script, world, port, *args, timeout=900):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(REPO, "tests", script), *args]
    r = subprocess.run(cmd, cwd=REPO, capture_output=True, text=True, timeout=timeout)
    print(r.stdout[-4000:])
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    return r.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 4, 8])
def test_tp_sharded_reference_stack_matches_unsharded(world):
    if world not in _worlds():
        pytest.skip(f"needs {world} GPUs")
    out = _torchrun("multi_gpu_tp.py", world, 29520 + world)
    assert '"ok": true' in out or '"skipped"' in out


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 4, 8])
def test_fused_oproj_allreduce_multi_gpu(world):
    if world not in _worlds():
        pytest.skip(f"needs {world} GPUs")
    _torchrun("multi_gpu_oproj.py", world, 29530 + world)
