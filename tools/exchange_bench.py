"""Exchange arithmetic (a2-a4) as blocking C-ABI calls, us/call (diagnostic): python tools/exchange_bench.py
sl_merge / sl_delta_snapshot / sl_apply_reply / sl_gossip_p2p / sl_train_tick at D = 1024 and 16384."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from serverless_learn_b200 import Replica  # noqa: E402


def timeit(fn, n=2000):
    for _ in range(50):
        fn()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    return (time.perf_counter() - t0) / n * 1e6


for D in (1024, 16384):
    with Replica(D, 64) as a, Replica(D, 64) as b:
        d = np.random.default_rng(0).normal(size=D)
        a.set_state(d, np.zeros(D)); b.set_state(-d, np.zeros(D))
        res = {
            "sl_merge": timeit(lambda: a.exchange_updates(d)),
            "sl_delta_snapshot": timeit(lambda: a.delta()),
            "sl_apply_reply": timeit(lambda: a.apply_reply(d)),
            "sl_gossip_p2p (whole round, one kernel)": timeit(lambda: a.gossip_p2p(b)),
            "sl_train_tick": timeit(lambda: a.train_tick()),
        }
        print(f"D={D}: " + "; ".join(f"{k} {v:.1f} us" for k, v in res.items()), flush=True)
