"""GPU: the role cores (WorkerCore / MasterCore / FileServerCore, include/sl_roles.h) run a
whole in-process job — registration, check-ups, file push, training ticks, gossip, exchange
with the master — and every replica is compared with an oracle simulation of the same
deterministic schedule (repair R8: one schedule instead of racing threads)."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_reference_job_in_process(orc, file0):
    from serverless_learn_b200 import roles
    F, B, NW = 1024, 4096, 3
    d, fs = roles.Directory(), roles.FileServer()
    master = roles.Master(F)
    ws = [roles.Worker(f"localhost:{50060 + k}", F, B) for k in range(NW)]
    for w in ws:
        d.add(w)
        assert master.register_birth(w.addr)           # src/worker.cc:249-252
    assert master.workers() == [w.addr for w in ws]    # registration order, append-only
    assert master.do_checkups(fs, d) == NW             # src/master.cc:240-266
    for w in ws:
        assert w.peers() == [x.addr for x in ws]       # full list, self included
    # push: every registered worker gets file 0 (src/master.cc:231-235)
    assert master.request_pushes(fs, d) == NW
    L = roles._lib.load()
    for w in ws:
        s, wsum = C.c_ulonglong(), C.c_ulonglong()
        assert L.sl_shard_checksum(w.handle, C.byref(s), C.byref(wsum)) == 0
        assert s.value == 12_749_417_848 and wsum.value == orc.weighted_sum(file0[0])
    # oracle replicas
    om = [np.zeros(F) for _ in range(NW)]
    oo = [np.zeros(F) for _ in range(NW)]
    mm, mo = np.zeros(F), np.zeros(F)
    # the gossip partner is rand() % N of THIS process: replay the same glibc stream
    libc = C.CDLL("libc.so.6")
    libc.srand(12345)
    draws = [libc.rand() for _ in range(6)]
    libc.srand(12345)
    it = iter(draws)
    for rnd in range(2):
        for k, w in enumerate(ws):
            for _ in range(k + 1):                     # simulate_training ticks, src/worker.cc:225-227
                assert w.train_tick()
                orc.train_tick(om[k])
        for k, w in enumerate(ws):                     # periodically_gossip, src/worker.cc:194-219
            j = w.gossip_once(d)
            assert j == next(it) % NW
            delta = orc.delta(om[k], oo[k])
            if j == k:                                 # a worker may draw itself (list includes it)
                reply = orc.merge_callee(om[k], oo[k], delta)
            else:
                reply = orc.merge_callee(om[j], oo[j], delta)
            orc.apply_reply(om[k], oo[k], reply)
        # R9: a worker also exchanges with the master (Master::ExchangeUpdates, proto:13)
        delta = orc.delta(om[0], oo[0])
        reply = master.exchange_updates(delta)
        assert np.array_equal(reply, orc.merge_callee(mm, mo, delta))
    for k, w in enumerate(ws):
        assert np.array_equal(w.model(), om[k]), k
    assert np.array_equal(master.model(), mm)
    # wire-level handler == plain handler
    raw = orc.update_encode(np.arange(F, dtype=np.float64))
    back = orc.update_decode(master.exchange_updates_wire(raw))
    assert np.array_equal(back, orc.merge_callee(mm, mo, np.arange(F, dtype=np.float64)))
    for w in ws:
        w.close()
    master.close()
    fs.close()
    d.close()


def test_gossip_without_peers_skips_the_round():
    from serverless_learn_b200 import roles
    d = roles.Directory()
    w = roles.Worker("a:1", 8, 4)
    assert w.gossip_once(d) == -1                      # R7 (reference: rand() % 0)
    w.check_up(["ghost:9"])
    assert w.gossip_once(d) == -1                      # unreachable partner
    w.close()
    d.close()


def test_worker_trains_on_pushed_synthetic_shard(orc):
    from serverless_learn_b200 import roles
    F, B, n = 64, 32, 500
    X, y = orc.synth(0, n, F)
    w = roles.Worker("a:1", F, B, eta=0.3)
    stream = np.concatenate([X.reshape(-1), y]).view(np.uint8)
    chunks = [stream[i:i + 1000] for i in range(0, stream.size, 1000)]   # rows straddle chunks
    assert w.receive_file(chunks)
    lg = w.train_steps(10)
    ref = np.zeros(F)
    lo = orc.lr_steps(X, y, B, 0, 10, ref, 0.3)
    assert np.max(np.abs(w.model() - ref)) / np.max(np.abs(ref)) <= 1e-5
    assert np.max(np.abs(lg - lo) / lo) <= 1e-5
    w.close()


def test_sharded_push_then_training(orc):
    # §8f n2: Push.file_num selects a shard; worker i gets file i+1 through DoPush/ReceiveFile
    # (1,000,000-byte chunks straddling rows), trains on it, matches the oracle on that shard
    from serverless_learn_b200 import roles
    F, B, n = 260, 100, 1500
    d, fs, master = roles.Directory(), roles.FileServer(), roles.Master(F)
    ws = [roles.Worker(f"w{k}:1", F, B, eta=0.2) for k in range(2)]
    for k, w in enumerate(ws):
        d.add(w)
        master.register_birth(w.addr)
        assert fs.add_synthetic_shard(k + 1, k * n, n, F)
    assert master.request_pushes_sharded(fs, d) == 2
    for k, w in enumerate(ws):
        lg = w.train_steps(12)
        X, y = orc.synth(k * n, n, F)
        ref = np.zeros(F)
        lo = orc.lr_steps(X, y, B, 0, 12, ref, 0.2)
        assert np.max(np.abs(w.model() - ref)) / np.max(np.abs(ref)) <= 1e-5
        assert np.max(np.abs(lg - lo) / lo) <= 1e-5
        w.close()
    master.close(); fs.close(); d.close()


def test_async_gossip_training_follows_the_oracle_schedule(orc):
    # §8f n3: the reference's actual algorithm — every worker trains on its own shard and, at
    # intervals, exchanges deltas with a rand()%N partner from its CheckUp peer list — run as one
    # deterministic schedule (R8) on the GPU replicas and in the oracle; also compared with the
    # synchronous (all-reduce) extension for convergence.
    from serverless_learn_b200 import roles
    F, B, n, NW, K, ROUNDS, eta = 64, 32, 640, 4, 5, 6, 0.3
    d, fs, master = roles.Directory(), roles.FileServer(), roles.Master(F)
    ws = [roles.Worker(f"w{k}:1", F, B, eta=eta) for k in range(NW)]
    shards = []
    for k, w in enumerate(ws):
        d.add(w)
        master.register_birth(w.addr)
        fs.add_synthetic_shard(k + 1, k * n, n, F)
        shards.append(orc.synth(k * n, n, F))
    assert master.request_pushes_sharded(fs, d) == NW and master.do_checkups(fs, d) == NW
    om = [np.zeros(F) for _ in range(NW)]
    oo = [np.zeros(F) for _ in range(NW)]
    libc = C.CDLL("libc.so.6")
    libc.srand(777)
    draws = [libc.rand() for _ in range(NW * ROUNDS)]
    libc.srand(777)
    it = iter(draws)
    gpu_loss, orc_loss = [], []
    for rnd in range(ROUNDS):
        for k, w in enumerate(ws):
            gpu_loss.append(w.train_steps(K))
            orc_loss.append(orc.lr_steps(shards[k][0], shards[k][1], B, rnd * K, K, om[k], eta))
        for k, w in enumerate(ws):
            j = w.gossip_once(d)
            assert j == next(it) % NW
            delta = orc.delta(om[k], oo[k])
            reply = orc.merge_callee(om[j], oo[j], delta)
            orc.apply_reply(om[k], oo[k], reply)
    for k, w in enumerate(ws):
        err = np.max(np.abs(w.model() - om[k])) / np.max(np.abs(om[k]))
        assert err <= 1e-5, (k, err)
    gl, ol = np.concatenate(gpu_loss), np.concatenate(orc_loss)
    assert np.max(np.abs(gl - ol) / ol) <= 1e-5
    # convergence: gossip keeps the replicas learning the same rule as synchronous averaging
    wsync = np.zeros(F)
    ls = orc.lr_steps_sync([s[0] for s in shards], [s[1] for s in shards], B, 0, K * ROUNDS, wsync, eta)
    last_async = np.mean([orc_loss[-NW + k][-1] for k in range(NW)])
    assert last_async < 0.75 * np.log(2) and ls[-1] < 0.75 * np.log(2)
    for w in ws:
        w.close()
    master.close(); fs.close(); d.close()


def test_elastic_join_and_dead_worker_removal(orc):
    # §8f n4 (beyond the reference, opt-in): a worker that registers mid-run gets a free shard and
    # the master's current model over the existing wire surface; a worker that stops answering
    # check-ups is dropped, its shard goes to the next joiner.  The arithmetic of every step is
    # checked against the oracle's merge (src/master.cc:95-114) and LR step.
    from serverless_learn_b200 import roles
    F, B, n, eta, lam = 64, 32, 320, 0.3, 0.5
    d, fs, master = roles.Directory(), roles.FileServer(), roles.Master(F, lam=lam)
    master.enable_elastic(2)
    for k in range(3):
        assert fs.add_synthetic_shard(k + 1, k * n, n, F)
    shards = {k + 1: orc.synth(k * n, n, F) for k in range(3)}
    ws = [roles.Worker(f"w{k}:1", F, B, eta=eta, lam=lam) for k in range(2)]
    for w in ws:
        d.add(w)
        master.register_birth(w.addr)
    assert master.admit_pending(fs, d) == 2 and master.admit_pending(fs, d) == 0   # pushed once
    assert [master.shard_of(w.addr) for w in ws] == [1, 2]
    assert master.do_checkups(fs, d) == 2
    # oracle replicas: workers + master
    om = [np.zeros(F) for _ in ws]
    oo = [np.zeros(F) for _ in ws]
    mm, mo = np.zeros(F), np.zeros(F)
    libc = C.CDLL("libc.so.6")
    libc.srand(4242)
    draws = [libc.rand() for _ in range(8)]
    libc.srand(4242)
    it = iter(draws)
    for k, w in enumerate(ws):
        w.train_steps(6)
        orc.lr_steps(*shards[k + 1], B, 0, 6, om[k], eta)
    for _ in range(3):   # the master follows the workers (the reference's dead initiator, revived)
        j = master.gossip_once(d)
        assert j == next(it) % 2
        reply = orc.merge_callee(om[j], oo[j], orc.delta(mm, mo))
        orc.apply_reply(mm, mo, reply)
    assert np.max(np.abs(mm)) > 0
    assert np.max(np.abs(master.model() - mm)) / np.max(np.abs(mm)) <= 1e-5   # trained weights: fp32 tolerance
    # ---- a third worker joins mid-run
    late = roles.Worker("w2:1", F, B, eta=eta, lam=lam)
    d.add(late)
    master.register_birth(late.addr)
    assert master.admit_pending(fs, d) == 1
    assert master.shard_of(late.addr) == 3
    assert np.array_equal(late.model(), master.model())          # exact: lambda = 0.5
    lg = late.train_steps(4)                                       # trains on shard 3 from that model
    ref = master.model().copy()
    lo = orc.lr_steps(*shards[3], B, 0, 4, ref, eta)
    assert np.max(np.abs(late.model() - ref)) / np.max(np.abs(ref)) <= 1e-5
    assert np.max(np.abs(lg - lo) / lo) <= 1e-5
    assert master.do_checkups(fs, d) == 3 and len(late.peers()) == 3
    # the seed is not re-broadcast as news: the joiner's next delta is only what it learned itself
    back = ws[0].exchange_updates(np.zeros(F))
    assert back is not None
    # ---- worker 1 dies: two silent check-ups drop it, the survivors' peer lists shrink
    d.remove(ws[1].addr)
    assert master.do_checkups(fs, d) == 2 and len(master.workers()) == 3    # first miss: still listed
    assert master.do_checkups(fs, d) == 2 and master.workers() == ["w0:1", "w2:1"]
    assert master.do_checkups(fs, d) == 2 and ws[0].peers() == ["w0:1", "w2:1"]
    assert master.shard_of(ws[1].addr) == 0
    # ---- the next joiner inherits the freed shard 2
    heir = roles.Worker("w3:1", F, B, eta=eta, lam=lam)
    d.add(heir)
    master.register_birth(heir.addr)
    assert master.admit_pending(fs, d) == 1 and master.shard_of(heir.addr) == 2
    lg = heir.train_steps(3)
    ref = master.model().copy()
    lo = orc.lr_steps(*shards[2], B, 0, 3, ref, eta)
    assert np.max(np.abs(heir.model() - ref)) / np.max(np.abs(ref)) <= 1e-5
    for w in ws + [late, heir]:
        w.close()
    master.close(); fs.close(); d.close()


def test_elastic_calls_are_off_by_default():
    from serverless_learn_b200 import roles
    d, fs, master = roles.Directory(), roles.FileServer(), roles.Master(16)
    master.register_birth("ghost:1")
    assert master.admit_pending(fs, d) == -1          # not enabled
    for _ in range(5):                                # reference behaviour: never removed
        assert master.do_checkups(fs, d) == 0
    assert master.workers() == ["ghost:1"]
    master.close(); fs.close(); d.close()


def test_described_shard_through_the_file_server(orc):
    # §8f n2: the file server writes the 64-byte header, the worker's ingest checks it
    from serverless_learn_b200 import roles
    F, B, n = 260, 100, 900
    d, fs = roles.Directory(), roles.FileServer()
    w = roles.Worker("w0:1", F, B, eta=0.2)
    d.add(w)
    assert fs.add_described_shard(1, 5000, n, F)
    assert fs.file_length_of(1) == 64 + n * (F + 1) * 4
    assert fs.do_push(d, w.addr, 1) == 1
    X, y = orc.synth(5000, n, F)
    ref = np.zeros(F)
    orc.lr_steps(X, y, B, 0, 7, ref, 0.2)
    w.train_steps(7)
    assert np.max(np.abs(w.model() - ref)) / np.max(np.abs(ref)) <= 1e-5
    # a shard written for another width is refused: the push reports ok=false (ReceiveFileAck.ok)
    assert fs.add_described_shard(2, 0, n, F // 2)
    assert fs.do_push(d, w.addr, 2) == 0
    w.train_steps(1)          # the previous shard is still trainable
    w.close(); fs.close(); d.close()


def test_config2_four_workers_on_one_gpu(orc):
    # BASELINE config 2 in small: master + 4 workers + file server in one process, every worker gets
    # its own shard through DoPush/ReceiveFile, then the four train synchronously as ONE kernel
    from serverless_learn_b200 import roles
    F, B, n, W, steps = 128, 64, 512, 4, 10
    d, fs, master = roles.Directory(), roles.FileServer(), roles.Master(F)
    ws = [roles.Worker(f"w{k}:1", F, B, eta=0.2) for k in range(W)]
    for k, w in enumerate(ws):
        d.add(w)
        master.register_birth(w.addr)
        assert fs.add_described_shard(k + 1, k * n, n, F)
    assert master.request_pushes_sharded(fs, d) == W
    lg = roles.train_steps_group(ws, steps)
    shards = [orc.synth(k * n, n, F) for k in range(W)]
    ref = np.zeros(F)
    lo = orc.lr_steps_sync([s[0] for s in shards], [s[1] for s in shards], B, 0, steps, ref, 0.2)
    for w in ws:
        assert np.max(np.abs(w.model() - ref)) / np.max(np.abs(ref)) <= 1e-5
    assert np.max(np.abs(lg - lo) / lo) <= 1e-5
    for w in ws:
        w.close()
    master.close(); fs.close(); d.close()


def test_workers_local_steps_side_by_side(orc):
    # the W training threads' bodies (src/worker.cc:221-231) as one call: every worker takes local steps
    # on its own shard, then a gossip round couples them — all against the oracle
    from serverless_learn_b200 import roles
    F, B, n, W, K = 128, 64, 512, 4, 6
    d, fs, master = roles.Directory(), roles.FileServer(), roles.Master(F)
    ws = [roles.Worker(f"w{k}:1", F, B, eta=0.2) for k in range(W)]
    for k, w in enumerate(ws):
        d.add(w)
        master.register_birth(w.addr)
        assert fs.add_synthetic_shard(k + 1, k * n, n, F)
    assert master.request_pushes_sharded(fs, d) == W and master.do_checkups(fs, d) == W
    shards = [orc.synth(k * n, n, F) for k in range(W)]
    om = [np.zeros(F) for _ in range(W)]
    oo = [np.zeros(F) for _ in range(W)]
    libc = C.CDLL("libc.so.6")
    libc.srand(99)
    draws = [libc.rand() for _ in range(2 * W)]
    libc.srand(99)
    it = iter(draws)
    for rnd in range(2):
        lg = roles.train_steps_multi(ws, K)
        for k in range(W):
            lo = orc.lr_steps(*shards[k], B, rnd * K, K, om[k], 0.2)
            assert np.max(np.abs(lg[k] - lo) / lo) <= 1e-5
        for k, w in enumerate(ws):
            j = w.gossip_once(d)
            assert j == next(it) % W
            reply = orc.merge_callee(om[j], oo[j], orc.delta(om[k], oo[k]))
            orc.apply_reply(om[k], oo[k], reply)
    for k, w in enumerate(ws):
        assert np.max(np.abs(w.model() - om[k])) / np.max(np.abs(om[k])) <= 1e-5
        w.close()
    master.close(); fs.close(); d.close()


@pytest.mark.parametrize("mode", ["async", "sync"])
def test_cpp_job_program_matches_the_oracle(orc, mode):
    # examples/sl_job.cc: the three roles in ONE C++ process (no Python in the loop) — file server, master
    # (RegisterBirth, sharded DoPush -> ReceiveFile with the self-describing header, CheckUp), W workers that
    # take local steps side by side and gossip with rand()%N partners (or train synchronously)
    import json
    import subprocess
    exe = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "serverless_learn_b200", "sl_job")
    if not os.path.exists(exe):
        pytest.skip("sl_job not built")
    W, n, F, B, R, K, eta = 3, 600, 128, 64, 4, 5, 0.2
    out = subprocess.run([exe, "--workers", str(W), "--rows", str(n), "--features", str(F), "--batch", str(B),
                          "--rounds", str(R), "--local-steps", str(K), "--eta", str(eta), "--mode", mode],
                         capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr
    got = json.loads(out.stdout.strip().splitlines()[-1])
    shards = [orc.synth(k * n, n, F) for k in range(W)]
    if mode == "sync":
        w = np.zeros(F)
        lo = orc.lr_steps_sync([s[0] for s in shards], [s[1] for s in shards], B, 0, R * K, w, eta)
        want = [lo[(r + 1) * K - 1] for r in range(R)]
        w00 = w[0]
    else:
        om = [np.zeros(F) for _ in range(W)]
        oo = [np.zeros(F) for _ in range(W)]
        libc = C.CDLL("libc.so.6")
        libc.srand(1)                                   # the program never seeds rand(): glibc's default
        want = []
        for r in range(R):
            last = []
            for k in range(W):
                last.append(orc.lr_steps(*shards[k], B, r * K, K, om[k], eta)[-1])
            want.append(float(np.mean(last)))
            for k in range(W):
                j = libc.rand() % W
                reply = orc.merge_callee(om[j], oo[j], orc.delta(om[k], oo[k]))
                orc.apply_reply(om[k], oo[k], reply)
        w00 = om[0][0]
    assert np.max(np.abs(np.array(got["mean_loss"]) - np.array(want)) / np.array(want)) <= 1e-5
    assert abs(got["w0_0"] - w00) <= 1e-5 * abs(w00)
