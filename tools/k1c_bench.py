"""Time the cluster kernel (K1c) over its variants, cluster counts and call lengths; optional per-step timeline.
    python tools/k1c_bench.py [--F 1024 --B 4096 --rows 1048576 --steps 2048] [--clv 1,2,3] [--clusters 0] [--timeline]
Timeline points (globaltimer ns): P0 step start, P5 weights taken, P1 compute done, P2 gradient sent (compute thread 0);
P6 partials landed, P7 red issued, P3 total seen, P4 weights broadcast (exchange thread 0)."""
import argparse
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from serverless_learn_b200 import Replica, SlError  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--F", type=int, default=1024)
ap.add_argument("--B", type=int, default=4096)
ap.add_argument("--rows", type=int, default=1 << 20)
ap.add_argument("--steps", default="2048")
ap.add_argument("--clv", default="1,2,3")
ap.add_argument("--clusters", default="0")
ap.add_argument("--ring", type=int, default=0)
ap.add_argument("--loss", type=int, default=1)
ap.add_argument("--path", type=int, default=6)
ap.add_argument("--delay", default="-1")
ap.add_argument("--coop", type=int, default=-1)
ap.add_argument("--graph", type=int, default=1)
ap.add_argument("--waitall", default="0")
ap.add_argument("--fused1", action="store_true", help="attach this replica to itself (1 rank): the cross-GPU stage runs against the own memory")
ap.add_argument("--gap", default="-1")
ap.add_argument("--timeline", action="store_true")
ap.add_argument("--span", type=int, default=0, help="steps of one call whose entry/exit stamps are compared with the event time")
a = ap.parse_args()
peak = 6535.7
algo = 4 * (a.F + 1) * a.B + 16 * a.F
names = ["P0 step start", "P1 compute done", "P2 gradient sent", "P3 total seen", "P4 weights broadcast",
         "P5 weights taken", "P6 partials landed", "P7 red issued", "P8 pass0 dots reduced", "P9 pass0 barrier+sum",
         "P10 pass0 sigmoid", "P11 pass0 gradient", "P12 pass1 rows loaded", "P13 pass1 dots reduced",
         "P14 pass1 barrier+sum", "P15 pass1 gradient"]
order = [0, 5, 8, 9, 10, 11, 12, 13, 14, 15, 1, 2, 6, 7, 3, 4]
with Replica(a.F, a.B) as r:
    r.shard_synth(a.rows)
    r.set_option("path", a.path)
    r.set_option("cl_coop", a.coop)
    r.set_option("cl_graph", a.graph)
    if a.fused1:
        r.peer_export(); r.peer_attach(1, 0, b"")
        r.train_steps = r.train_steps_sync
    for clv in [int(x) for x in a.clv.split(",")]:
        for ncl in [int(x) for x in a.clusters.split(",")]:
          for delay in [int(x) for x in a.delay.split(",")]:
           for gap in [int(x) for x in a.gap.split(",")]:
            r.set_option("poll_delay", delay); r.set_option("poll_gap", gap)

            for steps, wa in [(int(x), int(w)) for x in a.steps.split(",") for w in a.waitall.split(",")]:
                r.set_option("wait_all_owners", wa)
                r.set_option("cl_variant", clv); r.set_option("clusters", ncl); r.set_option("ring_slots", a.ring)
                r.set_option("debug_timeline", -1)
                try:
                    r.train_steps(64, want_loss=bool(a.loss))
                    ts = []
                    for _ in range(5):
                        r.train_steps(steps, want_loss=bool(a.loss))
                        ts.append(r.last_train_ms())
                    us = float(np.median(ts)) * 1e3 / steps
                    print(f"clv={clv} waitall={wa} delay={delay} gap={gap} grid={r.info('grid')} slots={r.info('ring_slots')} steps={steps} "
                          f"{us:8.3f} us/step (min {min(ts) * 1e3 / steps:.3f})  {algo / us / 1e3:8.1f} GB/s  "
                          f"frac={algo / us / 1e3 / peak:.3f}", flush=True)
                except SlError as e:
                    print(f"clv={clv} clusters={ncl}: {e}", flush=True)
                    continue
            if a.span:
                r.set_option("debug_timeline", 0)
                r.train_steps(a.span, want_loss=bool(a.loss))
                ms = r.last_train_ms()
                G = r.info("grid")
                buf = np.zeros(4 * G * 32, dtype=np.uint64)
                assert r._L.sl_debug_read(r._h, buf.ctypes.data_as(C.POINTER(C.c_ulonglong)), buf.size) == 0
                t = buf.reshape(4, G, 16, 2).astype(np.int64)
                ent, ext = t[0, :, 14, 1], t[0, :, 15, 1]
                e0 = ent.min()
                print(f"span of a {a.span}-step call: event time {ms * 1e3:.2f} us; kernel entry spread {ent.max() - e0} ns, "
                      f"exit {ext.min() - e0}..{ext.max() - e0} ns after the first entry")
                print(f"  cluster barrier passed {np.median(t[0, :, 9, 1]) - e0:.0f}  weights loaded {np.median(t[0, :, 10, 1]) - e0:.0f}  "
                      f"first rows landed {np.median(t[0, :, 11, 1]) - e0:.0f} (median ns after the first entry)")
                for s_ in range(min(4, a.span)):
                    print(f"  step {s_}: P0 {np.median(t[s_, :, 0, 1]) - e0:.0f}  P5 {np.median(t[s_, :, 5, 1]) - e0:.0f}  P1 {np.median(t[s_, :, 1, 1]) - e0:.0f}  "
                          f"P7 {np.median(t[s_, :, 7, 1]) - e0:.0f}  P3 {np.median(t[s_, :, 3, 1]) - e0:.0f}  P4 {np.median(t[s_, :, 4, 1]) - e0:.0f} (median ns after the first entry)")
            if a.timeline:
                r.set_option("debug_timeline", 10)
                r.train_steps(32, want_loss=bool(a.loss))
                G = r.info("grid")
                buf = np.zeros(4 * G * 32, dtype=np.uint64)
                assert r._L.sl_debug_read(r._h, buf.ctypes.data_as(C.POINTER(C.c_ulonglong)), buf.size) == 0
                t = buf.reshape(4, G, 16, 2).astype(np.int64)
                for s in range(1, 3):
                    g0 = t[s, :, 0, 1].min()
                    print(f"--- clv={clv} step {10 + s}: globaltimer ns after the earliest CTA's step start (min / median / max over CTAs)")
                    for p in order:
                        if not t[s, :, p, 1].any():
                            continue
                        v = t[s, :, p, 1] - g0
                        print(f"  {names[p]:22s} {v.min():7d} {int(np.median(v)):7d} {v.max():7d}")
                    ck = t[s, :, :, 0]
                    print("  SM-clock deltas (median over CTAs): " + ", ".join(
                        f"{names[b_][:2]}-{names[a_][:2]}={int(np.median(ck[:, b_] - ck[:, a_]))}"
                        for a_, b_ in [(0, 5), (5, 8), (8, 9), (9, 10), (10, 11), (11, 12), (12, 13), (13, 14), (14, 15), (5, 1), (1, 2), (6, 7), (7, 3), (3, 4)]))
                    for pp in (5, 1, 6):
                        late = np.argsort(-(t[s, :, pp, 1] - g0))[:6]
                        print(f"  latest {names[pp][:2]} CTAs:", late.tolist(), (t[s, late, pp, 1] - g0).tolist())
                    nxt = t[s + 1, :, 0, 1] - g0
                    print(f"  next step start        {nxt.min():7d} {int(np.median(nxt)):7d} {nxt.max():7d}")
