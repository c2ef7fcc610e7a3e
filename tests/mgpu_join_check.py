"""(Test infrastructure; lives under tests/ because it uses the oracle.)
A worker joins the synchronous group mid-run, the group dissolves, and forms again (run under torchrun, 2 ranks):
    alone (rank 0, K steps) -> join: rank 1 takes rank 0's model and step -> together -> sl_peer_detach on both ->
    alone again -> attach again -> together.
Every phase against the oracle's loops; the models bit-identical on both ranks whenever they train together.
The membership changes stand for src/master.cc:79-91 (birth) and :177-196 (a worker that stopped answering).
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tests/mgpu_join_check.py
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle  # noqa: E402  (checker)
from serverless_learn_b200 import Replica  # noqa: E402

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
assert world == 2
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
orc = oracle.Oracle()
F = int(os.environ.get("SL_F", 1024))
B = int(os.environ.get("SL_B", 4096))
n, eta = max(2 * B, (1 << 23) // F), 0.1
K = (10, 12, 7, 9)      # alone, together, alone, together


def gather_bytes(b: bytes, size: int) -> bytes:
    t = torch.frombuffer(bytearray(b), dtype=torch.uint8).cuda()
    out = [torch.zeros(size, dtype=torch.uint8, device="cuda") for _ in range(world)]
    dist.all_gather(out, t)
    return b"".join(bytes(o.cpu().numpy().tobytes()) for o in out)


def model_of_rank0(r) -> np.ndarray:
    t = torch.from_numpy(r.model().copy()).cuda()
    dist.broadcast(t, 0)
    return t.cpu().numpy()


def same_on_both(r) -> bool:
    allw = [torch.zeros(F, dtype=torch.float64, device="cuda") for _ in range(world)]
    dist.all_gather(allw, torch.from_numpy(r.model().copy()).cuda())
    return bool(torch.equal(allw[0], allw[1]))


def werr(a, b):
    return float(np.max(np.abs(a - b)) / np.max(np.abs(b)))


shards = [orc.synth(k * n, n, F) for k in range(world)]
Xs, ys = [s[0] for s in shards], [s[1] for s in shards]
w = np.zeros(F)
t = 0
errs = []
paths = []
with Replica(F, B, eta=eta, device=lr) as r:
    r.load_shard(*shards[rank])
    for phase, k in enumerate(K):
        if phase % 2 == 0:
            # rank 0 alone (never attached, or detached): the one-GPU kernels
            lo = orc.lr_steps(Xs[0], ys[0], B, t, k, w, eta)
            if rank == 0:
                lg = r.train_steps(k)
                errs.append((werr(r.model(), w), float(np.max(np.abs(lg - lo) / lo))))
                paths.append(r.info("path"))
            t += k
            dist.barrier()
        else:
            # join: the newcomer takes the group's state and its place in the minibatch schedule, then all attach
            m0 = model_of_rank0(r)
            if rank == 1:
                r.set_state(m0, None)
                r.set_option("step", t)
            r.peer_attach(world, rank, gather_bytes(r.peer_export(), 64))
            dist.barrier(); torch.cuda.synchronize()
            lo = orc.lr_steps_sync(Xs, ys, B, t, k, w, eta)
            r.train_steps_sync(k)
            assert same_on_both(r), f"phase {phase}: models differ between the ranks"
            errs.append((werr(r.model(), w), 0.0))
            paths.append(r.info("path"))
            t += k
            dist.barrier(); torch.cuda.synchronize()
            r.peer_detach()
            dist.barrier()
    assert r.info("step") == t or rank == 1
if rank == 0:
    print(f"join check F={F} B={B}: phases {K}, paths {paths}, (w_err, loss_err) per phase:",
          " ".join(f"({a:.1e},{b:.1e})" for a, b in errs), flush=True)
    assert all(a <= 1e-5 and b <= 1e-5 for a, b in errs), errs
    print("join check passed", flush=True)
dist.destroy_process_group()
