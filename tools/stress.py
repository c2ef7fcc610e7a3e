"""Long-run determinism / agreement stress (diagnostic): python tools/stress.py [steps multiplier]
Every case: two runs of the persistent kernel must be bit-identical (weights and every loss); the
simple three-launch path on the same data must agree within the fp32 tolerance."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from serverless_learn_b200 import Replica  # noqa: E402

CASES = [  # F, B, rows, steps, variant
    (1024, 4096, 1 << 16, 6000, 1), (1024, 4096, 1 << 16, 3000, 2), (1024, 4096, 1 << 16, 3000, 10),
    (1024, 1000, 50000, 6000, 1), (2048, 2048, 1 << 15, 3000, 5), (2048, 2048, 1 << 15, 3000, 24),
    (4096, 1024, 1 << 14, 3000, 22), (4096, 4096, 1 << 14, 1500, 15), (8192, 2048, 1 << 13, 1500, 14),
    (8192, 1024, 1 << 13, 1500, 21), (16384, 2048, 1 << 12, 800, 11), (16384, 512, 1 << 12, 1500, 26),
    (256, 4096, 1 << 17, 6000, 1), (1024, 148, 9999, 6000, 19),
]
SCALE = int(sys.argv[1]) if len(sys.argv) > 1 else 1
bad = 0
for F, B, rows, steps, variant in CASES:
    steps *= SCALE
    out = []
    t0 = time.time()
    for rep in range(2):
        with Replica(F, B, eta=0.05) as r:
            r.shard_synth(rows, row0=12345)
            r.set_option("path", 1)
            r.set_option("variant", variant)
            loss = r.train_steps(steps)
            out.append((r.model(), loss))
    same = np.array_equal(out[0][0], out[1][0]) and np.array_equal(out[0][1], out[1][1])
    ref_steps = min(steps, 400)
    with Replica(F, B, eta=0.05) as r:
        r.shard_synth(rows, row0=12345)
        r.set_option("path", 2)
        l2 = r.train_steps(ref_steps)
    with Replica(F, B, eta=0.05) as r:
        r.shard_synth(rows, row0=12345)
        r.set_option("path", 1)
        r.set_option("variant", variant)
        l1 = r.train_steps(ref_steps)
        w1 = r.model()
    el = float(np.max(np.abs(l1 - l2) / np.abs(l2)))
    ok = same and el <= 1e-5 and np.all(np.isfinite(out[0][1]))
    bad += 0 if ok else 1
    print(f"F={F:6d} B={B:5d} variant={variant:2d} steps={steps}: identical={same} loss_vs_simple={el:.2e} "
          f"{'ok' if ok else 'FAIL'} ({time.time() - t0:.1f}s)", flush=True)
print("stress:", "all ok" if bad == 0 else f"{bad} FAILED")
sys.exit(1 if bad else 0)
