"""Multi-GPU parity where the test runner can see it: one process per GPU (torch.distributed.run), the fused
peer-memory kernels of sl_train_steps_sync against the oracle's W-worker loop, models bit-identical on all
ranks.  Skipped on a box with one GPU.  The path stands for the reference's ExchangeUpdates between worker
processes (src/worker.cc:81-100 <-> :194-219), here as the synchronous extension.
"""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _run(world, env_extra, script="mgpu_sync_check.py", expect="model bit-identical on all ranks: True"):
    env = dict(os.environ, SL_BIG=str(1 << 16), **env_extra)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           os.path.join(ROOT, "tests", script)]
    p = subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=600)
    out = p.stdout + p.stderr
    assert p.returncode == 0, out[-3000:]
    assert expect in out, out[-3000:]
    return out


@pytest.mark.parametrize("F,B", [(1024, 4096), (256, 100), (4096, 512)])
def test_sync_training_across_gpus_matches_the_oracle(F, B):
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least two GPUs")
    world = 2 if n < 4 else 4
    out = _run(world, {"SL_F": str(F), "SL_B": str(B)})
    assert f"world={world}" in out


@pytest.mark.parametrize("F,B", [(1024, 4096), (4096, 512)])
def test_worker_joins_and_leaves_the_sync_group(F, B):
    # alone -> join (state + step handed over) -> together -> detach -> alone -> attach again -> together
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs at least two GPUs")
    _run(2, {"SL_F": str(F), "SL_B": str(B)}, script="mgpu_join_check.py", expect="join check passed")
