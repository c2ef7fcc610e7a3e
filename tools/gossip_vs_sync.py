"""The reference's asynchronous gossip (§8f n3) next to the synchronous extension, same data, one GPU
(diagnostic): python tools/gossip_vs_sync.py
W workers hold their own shards.  Async: every round each worker trains K local steps (all workers side by side in one kernel,
sl_train_steps_multi), then gossips once
with a rand()%N partner from its CheckUp peer list (one-kernel P2P round).  Sync: the W workers take the
same number of steps as one job (sl_train_steps_group).  Prints the mean training loss along the way."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from serverless_learn_b200 import roles  # noqa: E402

W, F, B, ROWS, K, ROUNDS, ETA = 4, 1024, 4096, 1 << 16, 16, 48, 0.1


def make():
    d, fs, master = roles.Directory(), roles.FileServer(), roles.Master(F)
    ws = [roles.Worker(f"w{k}:1", F, B, eta=ETA) for k in range(W)]
    for k, w in enumerate(ws):
        d.add(w)
        master.register_birth(w.addr)
        w.replica_shard_synth(ROWS, k * ROWS)
    master.do_checkups(fs, d)
    return d, fs, master, ws


def close(d, fs, master, ws):
    for w in ws:
        w.close()
    master.close(); fs.close(); d.close()


rows = []
d, fs, master, ws = make()
t0 = time.perf_counter()
for rnd in range(ROUNDS):
    losses = roles.train_steps_multi(ws, K)[:, -1]   # the four training threads' K steps, one kernel
    for w in ws:
        w.gossip_once(d)
    if (rnd + 1) % 8 == 0:
        rows.append(((rnd + 1) * K, float(np.mean(losses))))
t_async = time.perf_counter() - t0
spread = max(float(np.max(np.abs(w.model() - ws[0].model()))) for w in ws)
wnorm = float(np.max(np.abs(ws[0].model())))
close(d, fs, master, ws)

d, fs, master, ws = make()
t0 = time.perf_counter()
sync_rows = []
for rnd in range(ROUNDS):
    lg = roles.train_steps_group(ws, K)
    if (rnd + 1) % 8 == 0:
        sync_rows.append(((rnd + 1) * K, float(lg[-1])))
t_sync = time.perf_counter() - t0
close(d, fs, master, ws)

print(f"{W} workers x {ROWS} rows x {F}, B={B}, eta={ETA}; async: {K} local steps then one gossip round per worker")
print("| local steps per worker | async gossip: mean loss | synchronous: loss |")
print("|---:|---:|---:|")
for (s, la), (_, ls) in zip(rows, sync_rows):
    print(f"| {s} | {la:.5f} | {ls:.5f} |")
print(f"async: replicas differ by {spread:.3e} (max |w_i - w_0|, |w|max = {wnorm:.3f}); wall {t_async * 1e3:.1f} ms "
      f"for {ROUNDS} rounds; sync: wall {t_sync * 1e3:.1f} ms")
