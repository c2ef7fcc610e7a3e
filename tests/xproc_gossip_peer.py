"""Initiator side of the cross-process gossip test (tests/test_gpu_parity.py::test_gossip_across_processes):
a worker PROCESS of its own, as in the reference's deployment (src/worker.cc:233-258).  Talks to the test over
stdin/stdout, one line per message; does no arithmetic of its own."""
import os
import sys
import threading
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import serverless_learn_b200 as sl  # noqa: E402

n, dev, seed = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
rng = np.random.default_rng(seed)
am, ao = (rng.standard_normal(n) * 10.0 ** rng.integers(-6, 6, n) for _ in range(2))


def say(*parts):
    print(*parts, flush=True)


with sl.Replica(n, 1, device=dev, lam=0.5) as A:
    A.set_state(am, ao)
    say("HANDLES", A.state_export().hex())
    for line in sys.stdin:
        cmd = line.split()
        if not cmd:
            continue
        if cmd[0] == "ROUND":
            # src/worker.cc:194-219: the gossip thread owns the state for the round; a training thread that
            # arrives meanwhile waits (here: a tick, +1.0 on every element, started while the state is lent)
            with_tick = len(cmd) > 1 and cmd[1] == "tick"
            A.state_lend()
            done = threading.Event()
            t = None
            if with_tick:
                t = threading.Thread(target=lambda: (A.train_tick(), done.set()))
                t.start()
            say("LENT")
            assert sys.stdin.readline().split()[0] == "DONE"
            if with_tick:
                time.sleep(0.05)
                blocked = not done.is_set()
            A.state_reclaim()
            if with_tick:
                t.join()
                say("TICK", "blocked" if blocked else "ran-early")
            say("STATE", A.model().tobytes().hex(), A.old().tobytes().hex())
        elif cmd[0] == "QUIT":
            break
