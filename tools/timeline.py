"""Per-step timeline of the persistent kernel (diagnostic): python tools/timeline.py [variant] [B] [F] [rows]
K1:  P2 = gradient staged, P3 = counter bumped, P4 = total load issued.
K1w: P2 = partial stored + arrival bumped, P3 = all partials seen, P4 = owner published its slice."""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from serverless_learn_b200 import Replica  # noqa: E402

variant = int(sys.argv[1]) if len(sys.argv) > 1 else 2
F = int(sys.argv[3]) if len(sys.argv) > 3 else 1024
B = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
ROWS = int(sys.argv[4]) if len(sys.argv) > 4 else 1 << 20
with Replica(F, B) as r:
    r.shard_synth(ROWS)
    r.set_option("variant", variant)
    r.train_steps(8, want_loss=False)
    r.set_option("debug_timeline", 10)
    r.train_steps(32, want_loss=False)
    G = r.info("grid")
    buf = np.zeros(4 * G * 16, dtype=np.uint64)
    rc = r._L.sl_debug_read(r._h, buf.ctypes.data_as(C.POINTER(C.c_ulonglong)), buf.size)
    assert rc == 0
    t = buf.reshape(4, G, 8, 2).astype(np.int64)
names = ["P0 step start", "P1 compute done", "P2 gradient staged", "P3 counter bumped", "P4 total load issued", "P5 weights applied", "P6 exch saw gradient", "P7 reduce wait_group done"]
order = [0, 5, 1, 2, 6, 7, 3, 4]
for s in range(1, 3):
    g0 = t[s, :, 0, 1].min()
    print(f"--- step {10 + s}: globaltimer ns relative to earliest CTA step start (min / median / max over CTAs)")
    for p in order:
        if not t[s, :, p, 1].any():
            continue  # point not stamped by this kernel
        v = t[s, :, p, 1] - g0
        print(f"  {names[p]:24s} {v.min():7d} {int(np.median(v)):7d} {v.max():7d}")
    ck = t[s, :, :, 0]
    print("  SM-clock deltas (median over CTAs): " + ", ".join(
        f"{names[b][:2]}-{names[a][:2]}={int(np.median(ck[:, b] - ck[:, a]))}" for a, b in [(0, 5), (5, 1), (1, 2), (2, 3), (3, 4)]))
    nxt = t[s + 1, :, 0, 1] - g0
    print(f"  next step start          {nxt.min():7d} {int(np.median(nxt)):7d} {nxt.max():7d}")
    late = np.argsort(-(t[s, :, 2, 1] - g0))[:5]
    print("  latest P2 CTAs:", late.tolist(), (t[s, late, 2, 1] - g0).tolist())
