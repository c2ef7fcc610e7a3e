"""`Replica`: one model replica (a worker, or the master) on one GPU — a 1:1 wrapper of
the sl_* C ABI.  Method names follow the reference's handlers and loops:

    receive_file(chunks)      Worker::ReceiveFile            src/worker.cc:49-61
    exchange_updates(delta)   {Master,Worker}Impl::ExchangeUpdates  src/master.cc:95-114
    delta() / apply_reply()   periodically_gossip + WorkerStub::ExchangeUpdates
                                                             src/worker.cc:145-172,194-219
    train_tick()              simulate_training              src/worker.cc:221-231
    train_steps(n)            the minibatch step that takes simulate_training's slot
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib

_dp = C.POINTER(C.c_double)


class SlError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"sl error {code}: {msg}")
        self.code = code


def _d(a: np.ndarray):
    return a.ctypes.data_as(_dp)


class Replica:
    def __init__(self, F: int, B: int, eta: float = 0.1, lam: float = 0.5, device: int = 0,
                 D: int | None = None):
        self._L = _lib.load()
        self._h = C.c_void_p()
        self.F, self.B, self.D = F, B, F if D is None else D
        rc = self._L.sl_create(device, self.D, F, B, eta, lam, C.byref(self._h))
        if rc != 0:
            msg = self._L.sl_last_error(None)
            self._h = C.c_void_p()
            raise SlError(rc, msg.decode() if msg else "?")

    # ---- plumbing -----------------------------------------------------------------
    def _ck(self, rc: int):
        if rc != 0:
            msg = self._L.sl_last_error(self._h)
            raise SlError(rc, msg.decode() if msg else "?")

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._L.sl_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def set_option(self, key: str, value: int):
        self._ck(self._L.sl_set_option(self._h, key.encode(), int(value)))

    def info(self, key: str) -> int:
        return int(self._L.sl_get_info(self._h, key.encode()))

    # ---- ingest -------------------------------------------------------------------
    def shard_begin(self, size_hint: int = 0):
        self._ck(self._L.sl_shard_begin(self._h, size_hint))

    def shard_append(self, data) -> None:
        if isinstance(data, (bytes, bytearray, memoryview)):
            buf = np.frombuffer(data, dtype=np.uint8)
        else:
            buf = np.ascontiguousarray(data).view(np.uint8).reshape(-1)
        self._ck(self._L.sl_shard_append(self._h, buf.ctypes.data, buf.size))

    def shard_append_ptr(self, ptr: int, nbytes: int) -> None:
        self._ck(self._L.sl_shard_append(self._h, C.c_void_p(ptr), nbytes))

    def shard_end(self) -> int:
        n = C.c_longlong()
        self._ck(self._L.sl_shard_end(self._h, C.byref(n)))
        return n.value

    def receive_file(self, chunks) -> bool:
        """Worker::ReceiveFile: consume a stream of Chunk payloads; returns the ack's ok."""
        try:
            self.shard_begin(0)
            for ch in chunks:
                self.shard_append(ch)
            self.shard_end()
            return True
        except SlError:
            return False

    def load_shard(self, X: np.ndarray, y: np.ndarray) -> int:
        """Stream a shard in the wire layout (X bytes then y bytes)."""
        X = np.ascontiguousarray(X, dtype=np.float32)
        y = np.ascontiguousarray(y, dtype=np.float32)
        assert X.shape[0] == y.shape[0] and X.shape[1] == self.F
        self.shard_begin(X.nbytes + y.nbytes)
        self.shard_append(X)
        self.shard_append(y)
        return self.shard_end()

    def shard_synth(self, n_rows: int, row0: int = 0, seed: int = 0x5EED):
        self._ck(self._L.sl_shard_synth(self._h, n_rows, row0, seed))

    def shard_checksum(self):
        a, b = C.c_ulonglong(), C.c_ulonglong()
        self._ck(self._L.sl_shard_checksum(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def shard_read(self, offset: int, n: int) -> np.ndarray:
        out = np.empty(n, dtype=np.uint8)
        self._ck(self._L.sl_shard_read(self._h, offset, out.ctypes.data, n))
        return out

    def shard_bytes(self) -> int:
        return int(self._L.sl_shard_bytes(self._h))

    # ---- training -----------------------------------------------------------------
    def train_tick(self):
        self._ck(self._L.sl_train_tick(self._h))

    def train_steps(self, n_steps: int, want_loss: bool = True):
        loss = np.empty(max(n_steps, 1), dtype=np.float64)
        self._ck(self._L.sl_train_steps(self._h, n_steps, _d(loss) if want_loss else None))
        return loss[:n_steps] if want_loss else None

    def train_steps_sync(self, n_steps: int, want_loss: bool = True):
        loss = np.empty(max(n_steps, 1), dtype=np.float64)
        self._ck(self._L.sl_train_steps_sync(self._h, n_steps, _d(loss) if want_loss else None))
        return loss[:n_steps] if want_loss else None

    @staticmethod
    def train_steps_group(replicas, n_steps: int, want_loss: bool = True):
        """W worker replicas that share one GPU, trained as one job (sl_train_steps_group): every step
        each worker's minibatch gradient over its own shard is summed and all apply the mean."""
        first = replicas[0]
        arr = (type(first._h) * len(replicas))(*[r._h for r in replicas])
        loss = np.empty(max(n_steps, 1), dtype=np.float64)
        first._ck(first._L.sl_train_steps_group(arr, len(replicas), n_steps, _d(loss) if want_loss else None))
        return loss[:n_steps] if want_loss else None

    @staticmethod
    def train_steps_multi(replicas, n_steps: int, want_loss: bool = True):
        """W replicas sharing one GPU each take n_steps LOCAL steps as one kernel (sl_train_steps_multi);
        returns the losses as [W, n_steps]."""
        first = replicas[0]
        W = len(replicas)
        arr = (type(first._h) * W)(*[r._h for r in replicas])
        loss = np.empty(max(n_steps, 1) * W, dtype=np.float64)
        first._ck(first._L.sl_train_steps_multi(arr, W, n_steps, _d(loss) if want_loss else None))
        return loss[:n_steps * W].reshape(W, n_steps) if want_loss else None

    def last_train_ms(self) -> float:
        return float(self._L.sl_last_train_ms(self._h))

    # ---- exchange -----------------------------------------------------------------
    def exchange_updates(self, delta_in: np.ndarray) -> np.ndarray:
        """Callee side of ExchangeUpdates: returns the reply delta."""
        delta_in = np.ascontiguousarray(delta_in, dtype=np.float64)
        reply = np.empty(self.D, dtype=np.float64)
        self._ck(self._L.sl_merge(self._h, _d(delta_in), delta_in.size, _d(reply), reply.size))
        return reply

    def delta(self) -> np.ndarray:
        out = np.empty(self.D, dtype=np.float64)
        self._ck(self._L.sl_delta_snapshot(self._h, _d(out), out.size))
        return out

    def apply_reply(self, back: np.ndarray) -> None:
        back = np.ascontiguousarray(back, dtype=np.float64)
        self._ck(self._L.sl_apply_reply(self._h, _d(back), back.size))

    def model(self) -> np.ndarray:
        out = np.empty(self.D, dtype=np.float64)
        self._ck(self._L.sl_get_model(self._h, _d(out), out.size))
        return out

    def old(self) -> np.ndarray:
        out = np.empty(self.D, dtype=np.float64)
        self._ck(self._L.sl_get_old(self._h, _d(out), out.size))
        return out

    def set_state(self, model=None, old=None):
        m = None if model is None else np.ascontiguousarray(model, dtype=np.float64)
        o = None if old is None else np.ascontiguousarray(old, dtype=np.float64)
        n = m.size if m is not None else o.size
        self._ck(self._L.sl_set_state(self._h, _d(m) if m is not None else None,
                                      _d(o) if o is not None else None, n))

    # ---- same-box data parallel ----------------------------------------------------
    @staticmethod
    def comm_unique_id() -> bytes:
        L = _lib.load()
        buf = C.create_string_buffer(128)
        rc = L.sl_comm_unique_id(buf)
        if rc != 0:
            raise SlError(rc, (L.sl_last_error(None) or b"?").decode())
        return buf.raw

    def comm_init_rank(self, nranks: int, rank: int, uid: bytes):
        buf = C.create_string_buffer(uid, 128)
        self._ck(self._L.sl_comm_init_rank(self._h, nranks, rank, buf))

    def gossip_p2p(self, callee: "Replica") -> None:
        """One gossip round with `callee` (this replica initiates), entirely on the GPUs."""
        self._ck(self._L.sl_gossip_p2p(self._h, callee._h))

    # -- the same round between worker processes (include/sl_gpu.h: sl_state_*) --
    def state_export(self) -> bytes:
        buf = C.create_string_buffer(128)
        self._ck(self._L.sl_state_export(self._h, buf))
        return buf.raw

    def state_attach(self, handles: bytes, n: int, peer_lambda: float) -> int:
        peer = C.c_int(-1)
        self._ck(self._L.sl_state_attach(self._h, C.create_string_buffer(handles, 128), n, peer_lambda, C.byref(peer)))
        return peer.value

    def state_detach(self, peer: int) -> None:
        self._ck(self._L.sl_state_detach(self._h, peer))

    def state_lend(self) -> None:
        self._ck(self._L.sl_state_lend(self._h))

    def state_reclaim(self) -> None:
        self._ck(self._L.sl_state_reclaim(self._h))

    def gossip_p2p_remote(self, peer: int) -> None:
        """One gossip round initiated by the worker process whose state is attached as `peer`; runs here."""
        self._ck(self._L.sl_gossip_p2p_remote(self._h, peer))

    def peer_export(self) -> bytes:
        buf = C.create_string_buffer(64)
        self._ck(self._L.sl_peer_export(self._h, buf))
        return buf.raw

    def peer_attach(self, nranks: int, rank: int, handles: bytes):
        buf = C.create_string_buffer(handles, 64 * nranks) if handles else None
        self._ck(self._L.sl_peer_attach(self._h, nranks, rank, buf))

    def peer_detach(self):
        """Leave the synchronous group (every member does before a group of another size attaches)."""
        self._ck(self._L.sl_peer_detach(self._h))

    def peer_attach_symmetric(self, nranks: int, rank: int, ptrs, multicast_ptr: int, nbytes: int):
        """Symmetric allocation made by the caller (e.g. torch.distributed._symmetric_memory): ptrs[p] = rank p's
        buffer as mapped here, multicast_ptr = its multicast address or 0."""
        arr = (C.c_void_p * nranks)(*[C.c_void_p(int(p)) for p in ptrs])
        self._ck(self._L.sl_peer_attach_symmetric(self._h, nranks, rank, arr, C.c_void_p(int(multicast_ptr) or None), nbytes))

    def allreduce_avg(self):
        self._ck(self._L.sl_allreduce_avg(self._h))
