"""World-size-2 `gloo` test (CPU) of the N>1 host-side logic: rank -> shard assignment and
the synchronous averaging rule that `sl_train_steps_sync` implements with NCCL on the GPU
box (gradient summed over ranks in fp64, w -= eta/(B*nranks) * sum).  The per-rank gradient
comes from the oracle, the all-reduce from torch.distributed; the result must equal the
oracle's own multi-worker sync loop."""
import os
import socket
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)

WORKER = r'''
import os, sys
sys.path.insert(0, os.environ["SL_ROOT"])
import numpy as np, torch, torch.distributed as dist
import oracle, bench
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
orc = oracle.Oracle()
F, B, n, steps, eta = 32, 16, 96, 7, 0.25
row0 = bench.shard_row0(rank, n)              # same assignment bench.py uses per GPU
X, y = orc.synth(row0, n, F)
w = np.zeros(F)
losses = []
for s in range(steps):
    g, loss = orc.lr_grad(X, y, B, s, w)
    t = torch.from_numpy(np.concatenate([g * B, [loss * B]]))   # sums, as the GPU path all-reduces
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    t = t.numpy()
    w = w - eta / (B * world) * t[:F]
    losses.append(t[F] / (B * world))
if rank == 0:
    Xs, ys = zip(*[orc.synth(bench.shard_row0(k, n), n, F) for k in range(world)])
    wr = np.zeros(F)
    lr = orc.lr_steps_sync(list(Xs), list(ys), B, 0, steps, wr, eta)
    assert np.allclose(w, wr, rtol=0, atol=1e-14), np.abs(w - wr).max()
    assert np.allclose(losses, lr, rtol=1e-13)
    print("GLOO_OK", world)
dist.destroy_process_group()
'''


def test_sync_rule_world_size_2(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    env = dict(os.environ, SL_ROOT=ROOT)
    out = subprocess.run(
        [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
         "--master-addr", "127.0.0.1", "--master-port", str(port), str(script)],
        capture_output=True, text=True, env=env, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    assert "GLOO_OK 2" in out.stdout


def test_shard_assignment_is_disjoint_and_ordered():
    sys.path.insert(0, ROOT)
    import bench
    n = bench.ROWS_PER_GPU
    starts = [bench.shard_row0(k, n) for k in range(8)]
    assert starts == [k * n for k in range(8)]
