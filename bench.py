#!/usr/bin/env python
"""bench.py — training samples/s of the data-parallel hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

One "step" = one minibatch step of the build-defined logistic regression (the slot of
the reference's simulate_training, src/worker.cc:221-231) over B = 4096 rows of
F = 1024 fp32 features per GPU — BASELINE.json config 3 at N = 1, config 4 at N > 1
(one worker per GPU, gradients averaged every step).  Synthetic counter-based data
(SURVEY.md §8d), generated on the device; the 4 GiB shard per GPU is far larger than
L2 (126 MB), so consecutive steps never re-read cached rows.

Prints ONE JSON line (rank 0).  Keys beyond the base contract:
  roofline      achieved = algorithmic bytes per launch / CUDA-event time of the launch
  cpu_baseline  the oracle's LR step (a port: the reference has no gradient step) on
                one host thread — the reference runs one training thread per worker
  e2e           the same metric through the C ABI with HOST buffers: per step one
                minibatch is streamed from pinned host memory (sl_shard_*), one step is
                taken and its loss is read back
  repeats       the timed K-step call is repeated (barrier + synchronize around each) and
                the MEDIAN device time is reported; ms_per_step_all lists them all
  exchange_us   the update-exchange calls (merge / delta / apply / one-kernel gossip round)
                as blocking C-ABI calls, microseconds per call at D = 1024 and 16384
  N > 1         ranks_bit_identical (a hash of every rank's fp64 model after the warm-up)
                and sync_parity_err (first steps against the oracle's W-worker loop)
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

F = 1024
B = 4096
ROWS_PER_GPU = 1 << 20
ETA = 0.1
METRIC = "training samples/sec"
UNIT = "samples/s"


def shard_row0(rank: int, rows_per_gpu: int = ROWS_PER_GPU) -> int:
    """Worker k <-> GPU k owns rows [k*rows_per_gpu, (k+1)*rows_per_gpu) of the synthetic matrix."""
    return rank * rows_per_gpu


def algorithmic_bytes_per_step(f=F, b=B) -> int:
    # 4*(F+1) per sample (row + label) + 16*F per step (fp64 w read+write)
    return 4 * (f + 1) * b + 16 * f


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic_per_step():
    """DRAM bytes per step from the committed ncu capture (a citation of that run, not a measurement of this one)."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    try:
        with open(p) as fh:
            d = json.load(fh)
        return float(d["dram_bytes_per_step"]), "profiles/ncu_traffic.json: " + d.get("source", "ncu capture, N=1")
    except Exception:
        return None, None


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.samples, self.reasons = index, False, [], set()
        self.max_mhz = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.dev = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.dev, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def sample(self):
        nv = self.nv
        self.samples.append(nv.nvmlDeviceGetClockInfo(self.dev, nv.NVML_CLOCK_SM))
        r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.dev) if hasattr(
            nv, "nvmlDeviceGetCurrentClocksEventReasons") else nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.dev)
        names = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown",
                 0x40: "hw_thermal_slowdown", 0x80: "hw_power_brake_slowdown"}
        for bit, n in names.items():
            if r & bit:
                self.reasons.add(n)

    def run(self):
        if self.nv is None:
            return
        while not self.stop_flag:
            try:
                self.sample()
            except Exception:
                break
            time.sleep(0.002)

    def result(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": []}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


# --------------------------------------------------------------------------------------
def cpu_baseline_port(seconds: float = 8.0):
    """Oracle LR step on ONE host thread (reference structure: one training thread per
    worker, src/worker.cc:246), on a bounded sample of the config-3 workload."""
    import oracle
    orc = oracle.Oracle()
    n = 1 << 15                       # 32768 rows x 1024 = 128 MiB: larger than host LLC
    X, y = orc.synth(0, n, F)
    w = np.zeros(F)
    t0 = time.perf_counter()
    orc.lr_steps(X, y, B, 0, 8, w, ETA)
    per = (time.perf_counter() - t0) / 8
    steps = int(max(8, min(4000, seconds / per)))
    w[:] = 0
    t0 = time.perf_counter()
    orc.lr_steps(X, y, B, 0, steps, w, ETA)
    dt = time.perf_counter() - t0
    return {"value": steps * B / dt, "unit": UNIT, "cores": 1, "kind": "port",
            "sample": f"{steps} steps of B={B} over a {n}x{F} fp32 shard, 1 thread, "
                      f"{dt:.1f} s; oracle/sl_oracle_lr.c (the reference has no gradient step)",
            "host_cores_online": orc.max_threads()}


def run_reference(args, rank, world):
    """--impl reference: the CPU restatement with every host thread it can use — one
    worker (thread) per core, each on its own shard slice, the reference's own
    parallel structure (N worker processes, one training thread each)."""
    if rank != 0:
        return
    import oracle
    orc = oracle.Oracle()
    W = orc.max_threads()
    n = 1 << 14                                     # 64 MiB of rows per worker
    shards = [orc.synth(k * n, n, F) for k in range(W)]
    Xs, ys = [s[0] for s in shards], [s[1] for s in shards]
    ws = [np.zeros(F) for _ in range(W)]
    # a "step" here = one minibatch on every worker; bound the run to about a minute
    t0 = time.perf_counter()
    orc.lr_bench_workers(Xs, ys, B, 2, ws, ETA)
    per = (time.perf_counter() - t0) / 2
    budget = 60.0
    steps = int(max(1, min(args.steps, budget / max(per, 1e-9))))
    warm = int(max(1, min(args.warmup, 10.0 / max(per, 1e-9))))
    orc.lr_bench_workers(Xs, ys, B, warm, ws, ETA)
    dt = orc.lr_bench_workers(Xs, ys, B, steps, ws, ETA)
    value = W * steps * B / dt
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": warm, "ms_per_step": 1e3 * dt / steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "impl": "reference",
        "config": {"workload": f"config 3 shape: F={F}, B={B} per worker, {W} CPU workers x "
                               f"{n} rows (bounded sample)", "requested_steps": args.steps},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": W, "kind": "port",
                         "sample": f"{steps} steps x {W} workers x B={B}; oracle port (the "
                                   "reference's hot path does not compile and has no gradient "
                                   "step, see DESIGN.md)"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# --------------------------------------------------------------------------------------
def run_b200(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist

    from serverless_learn_b200 import Replica

    torch.cuda.set_device(local_rank)
    multi = world > 1
    numa_note = None
    if multi:
        # every rank's pinned host buffers (first touch) and its copy-issuing thread on the CPUs next to its GPU:
        # eight ranks streaming from one socket's memory halve each other's host-to-device rate
        try:
            import pynvml
            pynvml.nvmlInit()
            hnd = pynvml.nvmlDeviceGetHandleByIndex(local_rank)
            words = pynvml.nvmlDeviceGetCpuAffinity(hnd, (os.cpu_count() + 63) // 64)
            cpus = {64 * i + b for i, w in enumerate(words) for b in range(64) if (w >> b) & 1}
            cpus &= os.sched_getaffinity(0)
            if cpus:
                os.sched_setaffinity(0, cpus)
                numa_note = f"rank bound to the {len(cpus)} CPUs nearest its GPU"
        except Exception as ex:
            numa_note = f"no CPU binding ({ex!r})"
    if multi:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        torch.cuda.synchronize()
        if multi:
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if not multi:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    symm_keep = []

    def attach_peers(rep, width=F):
        """Exchange the 64-byte IPC handles of the replicas' inboxes: the fused kernels then sum gradients
        across GPUs through peer memory (NVLink), no per-step collective launch.  For narrow rows (cluster
        kernel) the all-rank words additionally go into a symmetric allocation with a multicast address where
        torch can make one: a GPU's total then reaches every rank with one multimem.st through the NVSwitch."""
        mine = torch.frombuffer(bytearray(rep.peer_export()), dtype=torch.uint8).cuda()
        allh = [torch.zeros(64, dtype=torch.uint8, device="cuda") for _ in range(world)]
        dist.all_gather(allh, mine)
        rep.peer_attach(world, rank, b"".join(bytes(t.cpu().numpy().tobytes()) for t in allh))
        if width <= 1024 and width % 16 == 0 and os.environ.get("SL_BENCH_MULTICAST", "1") != "0":
            try:
                import torch.distributed._symmetric_memory as symm
                nb = 128 * width
                tbuf = symm.empty(nb, dtype=torch.uint8, device=torch.device("cuda", local_rank))
                tbuf.zero_()
                hdl = symm.rendezvous(tbuf, dist.group.WORLD)
                torch.cuda.synchronize()
                dist.barrier()
                rep.peer_attach_symmetric(world, rank, [int(p) for p in hdl.buffer_ptrs], int(hdl.multicast_ptr or 0), nb)
                symm_keep.append((tbuf, hdl))
            except Exception as ex:  # no symmetric memory here: the IPC inbox (plain stores) stays
                print(f"bench.py: symmetric memory unavailable ({ex!r}); using per-peer stores", file=sys.stderr)

    r = Replica(F, B, eta=ETA, device=local_rank)
    r.shard_synth(ROWS_PER_GPU, row0=shard_row0(rank))
    if multi:
        uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            uid.copy_(torch.frombuffer(bytearray(Replica.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(uid, 0)
        r.comm_init_rank(world, rank, bytes(uid.cpu().numpy().tobytes()))
        attach_peers(r)
    step_fn = r.train_steps_sync if multi else r.train_steps

    # ---- N > 1: parity where the driver can see it (outside any timed region) ---------------
    sync_parity_err = None
    if multi:
        import oracle  # checker only
        psteps = 3
        lg = step_fn(psteps, want_loss=True)
        wg = torch.from_numpy(r.model()).cuda()
        allw = [torch.zeros(F, dtype=torch.float64, device="cuda") for _ in range(world)]
        dist.all_gather(allw, wg)
        if rank == 0:
            orc = oracle.Oracle()
            sh = [orc.synth(shard_row0(k), psteps * B, F) for k in range(world)]
            wref = np.zeros(F)
            lo = orc.lr_steps_sync([x[0] for x in sh], [x[1] for x in sh], B, 0, psteps, wref, ETA,
                                   threads=min(world, orc.max_threads()))
            ew = float(np.max(np.abs(allw[0].cpu().numpy() - wref)) / np.max(np.abs(wref)))
            el = float(np.max(np.abs(lg - lo) / lo))
            sync_parity_err = {"weights": ew, "loss": el, "steps": psteps,
                               "against": "oracle lr_steps_sync over the ranks' first minibatches"}
        r.set_state(np.zeros(F), np.zeros(F))
        r.set_option("step", 0)

    # ---- device-resident leg ("value") ------------------------------------------------
    step_fn(args.warmup, want_loss=True)
    ranks_bit_identical = None
    if multi:   # every rank must hold the same fp64 model, bit for bit
        mine = torch.from_numpy(r.model().view(np.int64).copy()).cuda()
        allm = [torch.zeros(F, dtype=torch.int64, device="cuda") for _ in range(world)]
        dist.all_gather(allm, mine)
        ranks_bit_identical = all(bool(torch.equal(allm[0], x)) for x in allm)
    sampler = ClockSampler(local_rank)
    barrier()
    l0 = r.info("launches")
    k0 = r.info("k1_launches")
    sampler.start()
    dev_all, wall_all = [], []
    for _ in range(args.repeats):            # EXACTLY K steps per repeat, barrier + synchronize around each
        barrier()
        t0 = time.perf_counter()
        step_fn(args.steps, want_loss=True)      # every step's loss is computed and read back
        torch.cuda.synchronize()
        wall_all.append(max_over_ranks(time.perf_counter() - t0))
        dev_all.append(max_over_ranks(r.last_train_ms() * 1e-3))
    sampler.stop_flag = True
    launches = (r.info("launches") - l0) // args.repeats
    path_used, variant_used, grid_used, slots_used = r.info("path"), r.info("variant"), r.info("grid"), r.info("ring_slots")
    multicast_used = r.info("multicast") == 1
    k1 = max(1, (r.info("k1_launches") - k0) // args.repeats) if path_used in (1, 4, 6, 7) else max(1, launches)
    barrier()
    sampler.join(timeout=1.0)
    dev_s = float(np.median(dev_all))
    wall_s = float(np.median(wall_all))
    dev_ms = dev_s * 1e3
    value = world * args.steps * B / dev_s
    r.close()                                 # its 4 GiB shard is not needed below
    barrier()

    # ---- end-to-end leg: host buffers through the C ABI ---------------------------------
    pool = 8
    host = [torch.empty(B * (F + 1), dtype=torch.float32).pin_memory() for _ in range(pool)]
    # host minibatches = the same synthetic rows, generated on the device by the product's
    # own generator and read back once, outside any timed region
    gen = Replica(F, B, eta=ETA, device=local_rank)
    for i, hb in enumerate(host):
        gen.shard_synth(B, row0=shard_row0(rank) + i * B)
        raw = gen.shard_read(0, gen.shard_bytes())
        hb.copy_(torch.from_numpy(raw.view(np.float32)))
    gen.close()
    e2e_steps = 256                           # long enough for the copy/compute pipeline to settle, whatever --steps is
    nbytes = B * (F + 1) * 4
    e2e = Replica(F, B, eta=ETA, device=local_rank)
    if multi:
        uid2 = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            uid2.copy_(torch.frombuffer(bytearray(Replica.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(uid2, 0)
        e2e.comm_init_rank(world, rank, bytes(uid2.cpu().numpy().tobytes()))
        attach_peers(e2e)
    e2e_step = e2e.train_steps_sync if multi else e2e.train_steps

    def ingest(i):
        hb = host[i % pool]
        e2e.shard_begin(nbytes)
        e2e.shard_append_ptr(hb.data_ptr(), nbytes)      # async DMA from pinned memory

    # every step: the NEXT minibatch's host->device copy is in flight while the current one is
    # trained on and its loss is read back; sl_shard_end waits for the copy and swaps it in
    ingest(0)
    e2e.shard_end()
    for i in range(16):
        ingest(i + 1)
        e2e_step(1, want_loss=True)
        e2e.shard_end()
    barrier()
    t0 = time.perf_counter()
    loss = 0.0
    for i in range(e2e_steps):
        ingest(i + 1)
        loss = e2e_step(1, want_loss=True)[0]
        e2e.shard_end()
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    barrier()
    e2e_value = world * e2e_steps * B / e2e_s

    # ---- extra: resident-shard job through the same C ABI (what the reference's design does:
    # a file is pushed once, then trained on): ingest a 2^16-row shard from pinned host memory,
    # train EPOCHS epochs on it, read every loss back.  Everything is inside the timed region.
    res_rows, epochs = 1 << 16, 64
    res_bytes = res_rows * (F + 1) * 4
    big = torch.empty(res_rows * (F + 1), dtype=torch.float32).pin_memory()
    gen = Replica(F, B, eta=ETA, device=local_rank)
    gen.shard_synth(res_rows, row0=shard_row0(rank))
    big.copy_(torch.from_numpy(gen.shard_read(0, gen.shard_bytes()).view(np.float32)))
    gen.close()
    res_steps = epochs * res_rows // B
    barrier()
    t0 = time.perf_counter()
    e2e.shard_begin(res_bytes)
    e2e.shard_append_ptr(big.data_ptr(), res_bytes)
    e2e.shard_end()
    e2e_step(res_steps, want_loss=True)
    torch.cuda.synchronize()
    res_s = max_over_ranks(time.perf_counter() - t0)
    barrier()
    del big

    # ---- extra: config 5's shape (wide rows, F = B = 16384) on the same box, device-resident ----
    # not the headline workload: one more line of evidence for the wide-row kernel (K1w)
    wide = None
    try:
        WF, WB, wrows, wsteps = 16384, 16384, 1 << 20, 64      # 2^20 rows x 16384 fp32 = 64 GiB resident per GPU
        rw = Replica(WF, WB, eta=ETA, device=local_rank)
        rw.shard_synth(wrows, row0=rank * wrows)
        if multi:
            uid3 = torch.zeros(128, dtype=torch.uint8, device="cuda")
            if rank == 0:
                uid3.copy_(torch.frombuffer(bytearray(Replica.comm_unique_id()), dtype=torch.uint8))
            dist.broadcast(uid3, 0)
            rw.comm_init_rank(world, rank, bytes(uid3.cpu().numpy().tobytes()))
            attach_peers(rw, WF)
        wstep = rw.train_steps_sync if multi else rw.train_steps
        wstep(8, want_loss=True)
        barrier()
        wstep(wsteps, want_loss=True)
        torch.cuda.synchronize()
        w_s = max_over_ranks(rw.last_train_ms() * 1e-3)
        barrier()
        wide = {"workload": f"config 5: F={WF}, B={WB} per GPU, {wrows} rows resident per GPU (64 GiB), "
                            f"{world * wrows} rows over {world} GPU(s)",
                "value": world * wsteps * WB / w_s, "unit": UNIT, "ms_per_step": 1e3 * w_s / wsteps,
                "steps": wsteps, "variant": rw.info("variant"),
                "hbm_gbs_per_gpu": algorithmic_bytes_per_step(WF, WB) * wsteps / w_s / 1e9}
        rw.close()
    except Exception as ex:  # the headline line must not depend on this extra
        wide = {"error": repr(ex)}

    # ---- extra: config 2 — the 4-worker, 4-shard job on ONE B200 (one kernel over the four
    # resident shards, sl_train_steps_group) next to the same job on 4 host threads ------------
    c2 = None
    if not multi:
        try:
            W2, rows2, steps2 = 4, 1 << 18, 2048
            reps = [Replica(F, B, eta=ETA, device=local_rank) for _ in range(W2)]
            for k, rp in enumerate(reps):
                rp.shard_synth(rows2, row0=k * rows2)
            Replica.train_steps_group(reps, 64, want_loss=True)
            torch.cuda.synchronize()
            Replica.train_steps_group(reps, steps2, want_loss=True)
            c2_s = reps[0].last_train_ms() * 1e-3
            c2 = {"workload": f"config 2: {W2} workers x {rows2} rows x {F} fp32 (4 GiB) resident on one B200, "
                              f"B={B} per worker per step, gradients summed every step",
                  "value": W2 * B * steps2 / c2_s, "unit": UNIT, "ms_per_step": 1e3 * c2_s / steps2, "steps": steps2,
                  "variant": reps[0].info("variant"),
                  "hbm_gbs": (W2 * 4 * (F + 1) * B + 16 * F) * steps2 / c2_s / 1e9}
            # the same four workers in the reference's own structure: local steps only (gossip couples
            # them every few seconds), all four side by side, a quarter of the clusters each
            Replica.train_steps_multi(reps, 64, want_loss=True)
            torch.cuda.synchronize()
            Replica.train_steps_multi(reps, steps2, want_loss=True)
            m_s = reps[0].last_train_ms() * 1e-3
            c2["local_steps"] = {"value": W2 * B * steps2 / m_s, "unit": UNIT, "ms_per_step": 1e3 * m_s / steps2,
                                 "what": f"{W2} independent replicas, {steps2} local steps each, side by side "
                                         "(sl_train_steps_multi: one cluster-kernel launch per replica, a "
                                         "quarter of the clusters each)",
                                 "kernel_path": reps[0].info("path")}
            for rp in reps:
                rp.close()
            import oracle
            orc = oracle.Oracle()
            n2 = 1 << 14
            sh = [orc.synth(k * n2, n2, F) for k in range(W2)]
            wv = np.zeros(F)
            t0 = time.perf_counter()
            orc.lr_steps_sync([x[0] for x in sh], [x[1] for x in sh], B, 0, 4, wv, ETA, threads=W2)
            per = (time.perf_counter() - t0) / 4
            cs = int(max(4, min(2000, 6.0 / per)))
            t0 = time.perf_counter()
            orc.lr_steps_sync([x[0] for x in sh], [x[1] for x in sh], B, 0, cs, wv, ETA, threads=W2)
            dt = time.perf_counter() - t0
            c2["cpu_4_workers"] = {"value": W2 * B * cs / dt, "unit": UNIT, "cores": W2, "kind": "port",
                                   "sample": f"{cs} synchronous steps, {W2} worker threads x B={B} over {n2}-row shards, {dt:.1f} s"}
        except Exception as ex:  # the headline line must not depend on this extra
            c2 = {"error": repr(ex)}

    # ---- extra: the update-exchange calls as blocking C-ABI calls, us per call ----
    xus = None
    if rank == 0:
        try:
            xus = {}
            for D in (1024, 16384):
                with Replica(D, 64, device=local_rank) as xa, Replica(D, 64, device=local_rank) as xb:
                    dvec = np.random.default_rng(0).normal(size=D)
                    xa.set_state(dvec, np.zeros(D))
                    xb.set_state(-dvec, np.zeros(D))

                    def per_call(fn, n=300):
                        for _ in range(30):
                            fn()
                        t0 = time.perf_counter()
                        for _ in range(n):
                            fn()
                        return (time.perf_counter() - t0) / n * 1e6

                    xus[f"D={D}"] = {"merge": per_call(lambda: xa.exchange_updates(dvec)),
                                     "delta": per_call(lambda: xa.delta()),
                                     "apply": per_call(lambda: xa.apply_reply(dvec)),
                                     "gossip_p2p": per_call(lambda: xa.gossip_p2p(xb))}
            xus["what"] = "host wall time per blocking call, host buffers in and out"
        except Exception as ex:
            xus = {"error": repr(ex)}

    if rank == 0:
        peak, peak_src = measured_peak()
        if c2 and "hbm_gbs" in c2:
            c2["roofline_frac"] = c2["hbm_gbs"] / peak
        if wide and "hbm_gbs_per_gpu" in wide:
            wide["roofline_frac"] = wide["hbm_gbs_per_gpu"] / peak
        algo = algorithmic_bytes_per_step() * args.steps / k1
        per_launch_s = dev_ms * 1e-3 / k1
        achieved = algo / per_launch_s / 1e9
        traffic, traffic_src = ncu_traffic_per_step()
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * dev_s / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": ("config 3" if world == 1 else "config 4") +
                                   f": logistic regression F={F}, B={B} per GPU, "
                                   f"{ROWS_PER_GPU} rows x {F} fp32 resident per GPU",
                       "l2": "inputs larger than L2 (4 GiB shard per GPU, sequential minibatches)",
                       "parallelism": f"dp{world}" + (" (gradient summed across GPUs every step inside the persistent "
                                                      "kernel, through peer memory over NVLink)" if multi else ""),
                       "kernel_path": {1: "persistent", 2: "simple", 3: "simple+nccl",
                                       4: "persistent+peer-memory exchange", 6: "cluster",
                                       7: "cluster+peer-memory exchange"}.get(path_used, "?"),
                       "multicast": bool(multicast_used),
                       "variant": variant_used, "grid": grid_used, "ring_slots": slots_used},
            "wall_ms_per_step": 1e3 * wall_s / args.steps,
            "repeats": args.repeats, "ms_per_step_all": [1e3 * x / args.steps for x in dev_all],
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak,
                         "traffic": None if (traffic is None or multi) else traffic * args.steps / k1,
                         "traffic_source": None if (traffic is None or multi) else traffic_src,
                         "peak_source": peak_src,
                         "algorithmic_bytes_per_step": algorithmic_bytes_per_step(),
                         "kernel": {1: "lr_persist_kernel", 4: "lr_persist_kernel", 6: "lr_cluster_kernel",
                                    7: "lr_cluster_kernel"}.get(path_used, "lr_fwd+lr_grad+lr_finish"),
                         "launches": k1},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": nbytes,
                    "d2h_bytes_per_step": 8, "steps": e2e_steps, "last_loss": float(loss),
                    "pcie_gbs": world * e2e_steps * nbytes / e2e_s / 1e9 / world, "host_binding": numa_note,
                    "note": "every sample crosses PCIe once and is used once: 4,100 B per sample bounds this leg at "
                            "host-to-device bandwidth / 4100, whatever the kernel does"},
            "e2e_resident": {"value": world * res_steps * B / res_s, "unit": UNIT,
                             "what": f"ingest {res_rows} rows ({res_bytes} B) from pinned host memory through "
                                     f"sl_shard_*, then {epochs} epochs = {res_steps} steps with every loss read "
                                     "back, all timed", "h2d_bytes": res_bytes, "d2h_bytes": 8 * res_steps},
            "gpu_launches": int(launches),   # kernels of one timed K-step call
            "clocks": sampler.result(),
            "config5": wide,
            "config2": c2,
            "exchange_us": xus,
        }
        if multi:
            line["ranks_bit_identical"] = ranks_bit_identical
            line["sync_parity_err"] = sync_parity_err
        if world == 1:
            line["cpu_baseline"] = cpu_baseline_port()
        emit(line)
    e2e.close()
    if multi:
        dist.destroy_process_group()


_REAL_STDOUT = None


def emit(line: dict) -> None:
    """The ONE JSON line goes to the real stdout; everything else any library prints to fd 1
    (NCCL's version banner, for one) has been routed to stderr."""
    out = os.fdopen(os.dup(_REAL_STDOUT), "w") if _REAL_STDOUT is not None else sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2048)
    ap.add_argument("--warmup", type=int, default=64)
    ap.add_argument("--repeats", type=int, default=9)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    args.repeats = max(args.repeats, 1)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if world != args.gpus and world == 1 and args.gpus > 1:
        print(f"bench.py: --gpus {args.gpus} needs torchrun (WORLD_SIZE={world})", file=sys.stderr)
        sys.exit(2)
    run_b200(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
