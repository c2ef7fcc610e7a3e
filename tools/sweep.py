"""Time the minibatch-step kernel over variants / grids (device events inside the lib).
    python tools/sweep.py [--F 1024 --B 4096 --rows 1048576 --steps 2048] [--variants 1,2,3]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from serverless_learn_b200 import Replica, SlError  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--F", type=int, default=1024)
ap.add_argument("--B", type=int, default=4096)
ap.add_argument("--rows", type=int, default=1 << 20)
ap.add_argument("--steps", type=int, default=2048)
ap.add_argument("--variants", default="1,2,3")
ap.add_argument("--grids", default="0")
ap.add_argument("--paths", default="1")
ap.add_argument("--ring", type=int, default=0)
ap.add_argument("--krep", type=int, default=0)
ap.add_argument("--skip", type=int, default=0)
ap.add_argument("--chunk", type=int, default=0)
ap.add_argument("--loss", type=int, default=0)
ap.add_argument("--block", type=int, default=0)
a = ap.parse_args()
peak = 6538.3
algo = 4 * (a.F + 1) * a.B + 16 * a.F
with Replica(a.F, a.B) as r:
    r.shard_synth(a.rows)
    for path in [int(x) for x in a.paths.split(",")]:
        for v in [int(x) for x in a.variants.split(",")]:
            for g in [int(x) for x in a.grids.split(",")]:
                r.set_option("ring_slots", a.ring); r.set_option("krep", a.krep); r.set_option("block_rows", a.block); r.set_option("debug_skip_compute", a.skip); r.set_option("tma_chunk", a.chunk); r.set_option("path", path); r.set_option("variant", v); r.set_option("grid", g)
                try:
                    r.train_steps(64, want_loss=False)
                    best = 1e9
                    for _ in range(3):
                        r.train_steps(a.steps, want_loss=bool(a.loss))
                        best = min(best, r.last_train_ms())
                    us = best * 1e3 / a.steps
                    print(f"path={path} variant={v} grid={r.info('grid')} slots={r.info('ring_slots')} "
                          f"{us:8.3f} us/step  {algo / us / 1e3:8.1f} GB/s  frac={algo / us / 1e3 / peak:.3f}",
                          flush=True)
                except SlError as e:
                    print(f"path={path} variant={v} grid={g}: {e}", flush=True)
                if path == 2:
                    break
            if path == 2:
                break
