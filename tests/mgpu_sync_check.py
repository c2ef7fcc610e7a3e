"""(Test infrastructure; lives under tests/ because it uses the oracle.)
Multi-GPU parity + timing of sl_train_steps_sync (run under torchrun, one rank per GPU):
the fused peer-memory kernel and the NCCL path against the oracle's W-worker sync loop.
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tests/mgpu_sync_check.py
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
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
orc = oracle.Oracle()
F = int(os.environ.get("SL_F", 1024))
B = int(os.environ.get("SL_B", 4096))
n, steps, eta = max(2 * B, (1 << 24) // F), 24, 0.1
BIG = int(os.environ.get("SL_BIG", 1 << 20))


def gather_bytes(b: bytes, size: int) -> bytes:
    t = torch.frombuffer(bytearray(b), dtype=torch.uint8).cuda()
    out = [torch.zeros(size, dtype=torch.uint8, device="cuda") for _ in range(world)]
    dist.all_gather(out, t)
    return b"".join(bytes(o.cpu().numpy().tobytes()) for o in out)


def bcast_bytes(b: bytes, size: int) -> bytes:
    t = torch.zeros(size, dtype=torch.uint8, device="cuda")
    if rank == 0:
        t.copy_(torch.frombuffer(bytearray(b), dtype=torch.uint8))
    dist.broadcast(t, 0)
    return bytes(t.cpu().numpy().tobytes())


shards = [orc.synth(k * n, n, F) for k in range(world)]
w = np.zeros(F)
lo = orc.lr_steps_sync([s[0] for s in shards], [s[1] for s in shards], B, 0, steps, w, eta)
res = {}
keep = []
for mode in ("fused", "fused_mc", "nccl"):
    r = Replica(F, B, eta=eta, device=lr)
    r.load_shard(*shards[rank])
    r.comm_init_rank(world, rank, bcast_bytes(Replica.comm_unique_id() if rank == 0 else b"", 128))
    if mode == "fused":
        r.peer_attach(world, rank, gather_bytes(r.peer_export(), 64))
    elif mode == "fused_mc":
        # the all-rank words in a symmetric allocation with a multicast address (torch's plumbing): one
        # multimem.st per column through the NVSwitch instead of a store per peer.  Cluster kernel only.
        if F > 1024 or F % 16:
            r.close()
            continue
        try:
            import torch.distributed._symmetric_memory as symm
            nb = 128 * F
            tbuf = symm.empty(nb, dtype=torch.uint8, device=torch.device("cuda", lr))
            tbuf.zero_()
            hdl = symm.rendezvous(tbuf, dist.group.WORLD)
            mc = int(hdl.multicast_ptr or 0)
            torch.cuda.synchronize(); dist.barrier()
            r.peer_attach(world, rank, gather_bytes(r.peer_export(), 64))
            r.peer_attach_symmetric(world, rank, [int(p) for p in hdl.buffer_ptrs], mc, nb)
            keep.append((tbuf, hdl))
            if rank == 0:
                print(f"fused_mc: symmetric memory attached, multicast address {'yes' if mc else 'NO (plain stores)'}", flush=True)
        except Exception as ex:  # no symmetric memory on this box / torch build: nothing to compare
            if rank == 0:
                print("fused_mc skipped:", repr(ex), flush=True)
            r.close()
            continue
    else:
        r.set_option("sync_path", 1)
    for kv in filter(None, os.environ.get("SL_OPTS", "").split(",")):   # e.g. SL_OPTS=wait_all_owners=1
        r.set_option(kv.split("=")[0], int(kv.split("=")[1]))
    for key in ("xpoll_delay", "xpoll_gap", "poll_delay", "poll_gap"):
        if os.environ.get("SL_" + key.upper()):
            r.set_option(key, int(os.environ["SL_" + key.upper()]))
    dist.barrier(); torch.cuda.synchronize()
    lg = r.train_steps_sync(steps)
    ew = float(np.max(np.abs(r.model() - w)) / np.max(np.abs(w)))
    el = float(np.max(np.abs(lg - lo) / lo))
    # timing on the big shard
    r.shard_synth(BIG, row0=rank * BIG)
    r.train_steps_sync(64, want_loss=False)
    dist.barrier(); torch.cuda.synchronize()
    if mode.startswith("fused") and os.environ.get("SL_TIMELINE"):
        import ctypes as C
        r.set_option("debug_timeline", 100)
        r.train_steps_sync(256, want_loss=False)
        G = r.info("grid")
        cl = r.info("path") == 7
        W_ = 32 if cl else 16
        buf = np.zeros(4 * G * W_, dtype=np.uint64)
        assert r._L.sl_debug_read(r._h, buf.ctypes.data_as(C.POINTER(C.c_ulonglong)), buf.size) == 0
        t = buf.reshape(4, G, W_ // 2, 2).astype(np.int64)
        if rank == 0:
            if cl:
                names = {0: "P0 step start", 5: "P5 weights taken", 1: "P1 compute done", 2: "P2 gradient sent",
                         6: "P6 partials landed", 7: "P7 red issued", 3: "P3 local total seen", 8: "P8 all ranks' totals seen",
                         4: "P4 weights broadcast"}
                order = (0, 5, 1, 2, 6, 7, 3, 8, 4)
            else:
                names = dict(enumerate(["P0 step start", "P1 compute done", "P2 staged", "P3 counter1 bumped",
                                        "P4 total load issued", "P5 weights applied", "P6 owner sent", "P7 owner got peers"]))
                order = (0, 5, 1, 2, 3, 6, 7, 4)
            for s_ in (1, 2):
                g0 = t[s_, :, 0, 1].min()
                print(f"--- fused step {100 + s_} (rank 0), ns from earliest step start: min/median/max over CTAs")
                for pnt in order:
                    v = t[s_, :, pnt, 1] - g0
                    v = v[t[s_, :, pnt, 1] > 0]
                    if v.size:
                        print(f"  {names[pnt]:26s} {v.min():7d} {int(np.median(v)):7d} {v.max():7d}   (n={v.size})")
                print(f"  next step start            {int(np.median(t[s_ + 1, :, 0, 1] - g0)):7d}", flush=True)
    r.train_steps_sync(1024, want_loss=True)
    ms = torch.tensor([r.last_train_ms()], device="cuda"); dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    res[mode] = (ew, el, r.info("path"), float(ms.item()) * 1e3 / 1024, r.model())
    r.close()
allw = [torch.zeros(F, dtype=torch.float64, device="cuda") for _ in range(world)]
dist.all_gather(allw, torch.from_numpy(res["fused"][4]).cuda())
same = all(bool(torch.equal(allw[0], x)) for x in allw)
if rank == 0:
    for mode, (ew, el, path, us, _) in res.items():
        print(f"{mode:6s} world={world} path={path} w_err={ew:.2e} loss_err={el:.2e}  {us:.2f} us/step "
              f"-> {world * B / us:.1f} M samples/s", flush=True)
    print("fused: model bit-identical on all ranks:", same, flush=True)
    assert res["fused"][0] <= 1e-5 and res["fused"][1] <= 1e-5 and res["nccl"][0] <= 1e-5 and same
    if "fused_mc" in res:
        assert res["fused_mc"][0] <= 1e-5 and res["fused_mc"][1] <= 1e-5
        assert np.array_equal(res["fused_mc"][4], res["fused"][4]), "multicast and per-peer stores must give the same bits"
dist.destroy_process_group()
