"""us/step of the automatically chosen kernel per shape, cluster kernel (path 6) beside the older persistent
kernels (path 1):  python tools/shapes_bench.py [--steps 1024]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from serverless_learn_b200 import Replica, SlError  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=1024)
ap.add_argument("--shapes", default="1024x4096,1024x8192,1024x16384,1024x1024,1024x512,512x4096,256x4096,992x4096,2048x4096,4096x8192,16384x16384")
a = ap.parse_args()
peak = 6535.7
for sh in a.shapes.split(","):
    F, B = (int(x) for x in sh.split("x"))
    rows = max(4 * B, (1 << 32) // (4 * F))          # 4 GiB of rows: larger than L2
    out = []
    for path in (0, 6, 1):
        try:
            with Replica(F, B) as r:
                r.shard_synth(rows)
                r.set_option("path", path)
                r.train_steps(64, want_loss=True)
                best = 1e9
                for _ in range(3):
                    r.train_steps(a.steps, want_loss=True)
                    best = min(best, r.last_train_ms())
                us = best * 1e3 / a.steps
                algo = 4 * (F + 1) * B + 16 * F
                out.append(f"path={path}: used {r.info('path')}/{r.info('variant')} {us:8.2f} us  frac={algo / us / 1e3 / peak:.3f}")
        except SlError as e:
            out.append(f"path={path}: n/a")
    print(f"F={F:6d} B={B:6d}  " + "  |  ".join(out), flush=True)
