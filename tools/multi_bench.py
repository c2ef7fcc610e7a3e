"""sl_train_steps_multi timing (diagnostic): python tools/multi_bench.py [rows per replica] [steps]
W = 2, 3, 4, 8 independent replicas of F = 1024, B = 4096 on one GPU: one cluster-kernel launch per replica (default),
and on the older kernels ("path" = 1) interleaved on every SM and as grid partitions (multi_mode = 1)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from serverless_learn_b200 import Replica  # noqa: E402

F, B = 1024, 4096
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 18
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
for W in (2, 3, 4, 8):
    for mode in (-1, 0, 1):
        reps = [Replica(F, B, eta=0.1) for _ in range(W)]
        for k, r in enumerate(reps):
            r.shard_synth(rows, row0=k * rows)
        if mode >= 0:
            reps[0].set_option("path", 1)
            reps[0].set_option("multi_mode", mode)
        Replica.train_steps_multi(reps, 64, want_loss=False)
        Replica.train_steps_multi(reps, steps, want_loss=False)
        ms = reps[0].last_train_ms()
        print(f"W={W} {('clusters   ', 'interleaved', 'partitions ')[mode + 1]} path={reps[0].info('path')} ring={reps[0].info('ring_slots')}: "
              f"{1e3 * ms / steps:.3f} us per step of all {W} -> {W * B * steps / ms / 1e3:.1f} M samples/s, "
              f"{W * (4 * (F + 1) * B + 16 * F) * steps / ms / 1e6:.0f} GB/s", flush=True)
        for r in reps:
            r.close()
