"""ctypes binding of libsl_roles.so (include/sl_roles.h): the C++ role cores that mirror the
reference's master / worker / file_server without gRPC.  Tests and examples only."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import _lib

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libsl_roles.so")
_dp = C.POINTER(C.c_double)
_p = C.c_void_p

_SIGS = {
    "slr_update_encode": (C.c_long, [_dp, C.c_int, _p, C.c_long]),
    "slr_update_decode": (C.c_int, [_p, C.c_long, _dp, C.c_int]),
    "slr_directory_create": (C.c_int, [C.POINTER(_p)]),
    "slr_directory_destroy": (None, [_p]),
    "slr_directory_add": (C.c_int, [_p, C.c_char_p, _p]),
    "slr_file_server_create": (C.c_int, [C.POINTER(_p)]),
    "slr_file_server_destroy": (None, [_p]),
    "slr_file_server_file_length": (C.c_longlong, [_p]),
    "slr_file_server_chunk": (C.c_longlong, [_p, C.c_longlong, C.POINTER(_p)]),
    "slr_file_server_do_push": (C.c_int, [_p, _p, C.c_char_p, C.c_uint]),
    "slr_file_server_check_up": (C.c_int, [_p]),
    "slr_file_server_add_synthetic_shard": (C.c_int, [_p, C.c_uint, C.c_ulonglong, C.c_ulonglong, C.c_uint,
                                                    C.c_ulonglong]),
    "slr_file_server_add_described_shard": (C.c_int, [_p, C.c_uint, C.c_ulonglong, C.c_ulonglong, C.c_uint,
                                                    C.c_ulonglong]),
    "slr_file_server_file_length_of": (C.c_longlong, [_p, C.c_uint]),
    "slr_file_server_chunk_of": (C.c_longlong, [_p, C.c_uint, C.c_longlong, C.POINTER(_p)]),
    "slr_master_request_pushes_sharded": (C.c_int, [_p, _p, _p]),
    "slr_master_create": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.POINTER(_p)]),
    "slr_master_destroy": (None, [_p]),
    "slr_master_handle": (_p, [_p]),
    "slr_master_register_birth": (C.c_int, [_p, C.c_char_p]),
    "slr_master_worker_count": (C.c_int, [_p]),
    "slr_master_worker_addr": (C.c_char_p, [_p, C.c_int]),
    "slr_master_exchange_updates": (C.c_int, [_p, _dp, C.c_int, _dp, C.c_int]),
    "slr_master_exchange_updates_wire": (C.c_int, [_p, _p, C.c_long, _p, C.c_long, C.POINTER(C.c_long)]),
    "slr_master_request_pushes": (C.c_int, [_p, _p, _p]),
    "slr_master_do_checkups": (C.c_int, [_p, _p, _p]),
    "slr_workers_train_steps_group": (C.c_int, [C.POINTER(_p), C.c_int, C.c_int, _dp]),
    "slr_workers_train_steps_multi": (C.c_int, [C.POINTER(_p), C.c_int, C.c_int, _dp]),
    "slr_master_enable_elastic": (C.c_int, [_p, C.c_int]),
    "slr_master_admit_pending": (C.c_int, [_p, _p, _p]),
    "slr_master_gossip_once": (C.c_int, [_p, _p]),
    "slr_master_shard_of": (C.c_int, [_p, C.c_char_p]),
    "slr_directory_remove": (C.c_int, [_p, C.c_char_p]),
    "slr_worker_create": (C.c_int, [C.c_int, C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_double,
                                    C.c_double, C.POINTER(_p)]),
    "slr_worker_destroy": (None, [_p]),
    "slr_worker_handle": (_p, [_p]),
    "slr_worker_addr": (C.c_char_p, [_p]),
    "slr_worker_receive_begin": (C.c_int, [_p]),
    "slr_worker_receive_chunk": (C.c_int, [_p, _p, C.c_size_t]),
    "slr_worker_receive_end": (C.c_int, [_p]),
    "slr_worker_check_up": (C.c_int, [_p, C.POINTER(C.c_char_p), C.c_int]),
    "slr_worker_peer_count": (C.c_int, [_p]),
    "slr_worker_peer_addr": (C.c_char_p, [_p, C.c_int]),
    "slr_worker_exchange_updates": (C.c_int, [_p, _dp, C.c_int, _dp, C.c_int]),
    "slr_worker_gossip_once": (C.c_int, [_p, _p]),
    "slr_worker_train_tick": (C.c_int, [_p]),
    "slr_worker_train_steps": (C.c_int, [_p, C.c_int, _dp]),
}
_lib_roles = None


def load() -> C.CDLL:
    global _lib_roles
    if _lib_roles is None:
        _lib.load()  # libsl_roles.so links against libsl_gpu.so
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: run `python -m serverless_learn_b200.build`")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGS.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _lib_roles = lib
    return _lib_roles


def _d(a):
    return a.ctypes.data_as(_dp)


def _model_of(handle, D):
    L = _lib.load()
    out = np.empty(D)
    assert L.sl_get_model(handle, _d(out), D) == 0
    return out


class Directory:
    def __init__(self):
        self.L, self.p = load(), _p()
        assert self.L.slr_directory_create(C.byref(self.p)) == 0

    def add(self, worker: "Worker"):
        assert self.L.slr_directory_add(self.p, worker.addr.encode(), worker.p) == 0

    def remove(self, addr: str):
        """The worker's server is gone (n4): calls to `addr` fail from now on."""
        assert self.L.slr_directory_remove(self.p, addr.encode()) == 1

    def close(self):
        if self.p:
            self.L.slr_directory_destroy(self.p)
            self.p = _p()


class FileServer:
    def __init__(self):
        self.L, self.p = load(), _p()
        assert self.L.slr_file_server_create(C.byref(self.p)) == 0

    def file_length(self) -> int:
        return int(self.L.slr_file_server_file_length(self.p))

    def chunk(self, k: int) -> np.ndarray:
        ptr = _p()
        n = self.L.slr_file_server_chunk(self.p, k, C.byref(ptr))
        if n == 0:
            return np.zeros(0, np.uint8)
        return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_uint8)), shape=(n,))

    def add_described_shard(self, file_num: int, row0: int, n_rows: int, F: int, seed: int = 0x5EED) -> bool:
        """Like add_synthetic_shard, with the 64-byte self-describing header in front (n2)."""
        return bool(self.L.slr_file_server_add_described_shard(self.p, file_num, row0, n_rows, F, seed))

    def add_synthetic_shard(self, file_num: int, row0: int, n_rows: int, F: int, seed: int = 0x5EED) -> bool:
        return bool(self.L.slr_file_server_add_synthetic_shard(self.p, file_num, row0, n_rows, F, seed))

    def file_length_of(self, file_num: int) -> int:
        return int(self.L.slr_file_server_file_length_of(self.p, file_num))

    def chunk_of(self, file_num: int, k: int) -> np.ndarray:
        ptr = _p()
        n = self.L.slr_file_server_chunk_of(self.p, file_num, k, C.byref(ptr))
        if n <= 0:
            return np.zeros(0, np.uint8)
        return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_uint8)), shape=(n,))

    def do_push(self, d: Directory, addr: str, file_num: int = 0) -> int:
        return int(self.L.slr_file_server_do_push(self.p, d.p, addr.encode(), file_num))

    def close(self):
        if self.p:
            self.L.slr_file_server_destroy(self.p)
            self.p = _p()


class Master:
    def __init__(self, D: int, B: int = 1, eta: float = 0.1, lam: float = 0.5, device: int = 0):
        self.L, self.p, self.D = load(), _p(), D
        rc = self.L.slr_master_create(device, D, B, eta, lam, C.byref(self.p))
        if rc != 0:
            raise RuntimeError(f"slr_master_create failed ({rc}): "
                               f"{(_lib.load().sl_last_error(None) or b'').decode()}")

    def register_birth(self, addr: str) -> bool:
        return bool(self.L.slr_master_register_birth(self.p, addr.encode()))

    def workers(self):
        n = self.L.slr_master_worker_count(self.p)
        return [self.L.slr_master_worker_addr(self.p, i).decode() for i in range(n)]

    def exchange_updates(self, delta):
        delta = np.ascontiguousarray(delta, dtype=np.float64)
        reply = np.empty(self.D)
        ok = self.L.slr_master_exchange_updates(self.p, _d(delta), delta.size, _d(reply), self.D)
        return reply if ok else None

    def exchange_updates_wire(self, raw: bytes):
        out = C.create_string_buffer(8 * self.D + 16)
        n = C.c_long()
        ok = self.L.slr_master_exchange_updates_wire(self.p, raw, len(raw), out, len(out), C.byref(n))
        return out.raw[:n.value] if ok else None

    def request_pushes(self, fs: FileServer, d: Directory) -> int:
        return int(self.L.slr_master_request_pushes(self.p, fs.p, d.p))

    def request_pushes_sharded(self, fs: FileServer, d: Directory) -> int:
        return int(self.L.slr_master_request_pushes_sharded(self.p, fs.p, d.p))

    def do_checkups(self, fs: FileServer, d: Directory) -> int:
        return int(self.L.slr_master_do_checkups(self.p, fs.p, d.p))

    # ---- elastic membership (beyond the reference, off unless enabled) ----
    def enable_elastic(self, max_missed: int = 3):
        assert self.L.slr_master_enable_elastic(self.p, max_missed) == 1

    def admit_pending(self, fs: FileServer, d: Directory) -> int:
        return int(self.L.slr_master_admit_pending(self.p, fs.p, d.p))

    def gossip_once(self, d: Directory) -> int:
        return int(self.L.slr_master_gossip_once(self.p, d.p))

    def shard_of(self, addr: str) -> int:
        return int(self.L.slr_master_shard_of(self.p, addr.encode()))

    def model(self):
        return _model_of(self.L.slr_master_handle(self.p), self.D)

    def close(self):
        if self.p:
            self.L.slr_master_destroy(self.p)
            self.p = _p()


def train_steps_group(workers, n: int):
    """W workers sharing one GPU take n synchronous steps as one kernel (config 2); returns the losses."""
    L = load()
    arr = (_p * len(workers))(*[w.p for w in workers])
    loss = np.empty(max(n, 1))
    assert L.slr_workers_train_steps_group(arr, len(workers), n, _d(loss)) == 1
    return loss[:n]


def train_steps_multi(workers, n: int):
    """W workers sharing one GPU each take n LOCAL steps side by side in one kernel; losses [W, n]."""
    L = load()
    W = len(workers)
    arr = (_p * W)(*[w.p for w in workers])
    loss = np.empty(max(n, 1) * W)
    assert L.slr_workers_train_steps_multi(arr, W, n, _d(loss)) == 1
    return loss[:n * W].reshape(W, n)


class Worker:
    def __init__(self, addr: str, F: int, B: int, eta: float = 0.1, lam: float = 0.5, device: int = 0):
        self.L, self.p, self.addr, self.D = load(), _p(), addr, F
        rc = self.L.slr_worker_create(device, addr.encode(), F, F, B, eta, lam, C.byref(self.p))
        if rc != 0:
            raise RuntimeError(f"slr_worker_create failed ({rc}): "
                               f"{(_lib.load().sl_last_error(None) or b'').decode()}")

    @property
    def handle(self):
        return self.L.slr_worker_handle(self.p)

    def replica_shard_synth(self, n_rows: int, row0: int = 0, seed: int = 0x5EED) -> None:
        """Fill this worker's shard on the device with synthetic rows (benchmarks; sl_shard_synth)."""
        assert _lib.load().sl_shard_synth(self.handle, n_rows, row0, seed) == 0

    def receive_file(self, chunks) -> bool:
        ok = bool(self.L.slr_worker_receive_begin(self.p))
        for ch in chunks:
            buf = np.ascontiguousarray(ch).view(np.uint8).reshape(-1)
            ok = bool(self.L.slr_worker_receive_chunk(self.p, buf.ctypes.data, buf.size)) and ok
        return bool(self.L.slr_worker_receive_end(self.p)) and ok

    def check_up(self, peers):
        arr = (C.c_char_p * len(peers))(*[p.encode() for p in peers])
        return bool(self.L.slr_worker_check_up(self.p, arr, len(peers)))

    def peers(self):
        n = self.L.slr_worker_peer_count(self.p)
        return [self.L.slr_worker_peer_addr(self.p, i).decode() for i in range(n)]

    def exchange_updates(self, delta):
        delta = np.ascontiguousarray(delta, dtype=np.float64)
        reply = np.empty(self.D)
        ok = self.L.slr_worker_exchange_updates(self.p, _d(delta), delta.size, _d(reply), self.D)
        return reply if ok else None

    def gossip_once(self, d: Directory) -> int:
        return int(self.L.slr_worker_gossip_once(self.p, d.p))

    def train_tick(self) -> bool:
        return bool(self.L.slr_worker_train_tick(self.p))

    def train_steps(self, n: int):
        loss = np.empty(max(n, 1))
        ok = self.L.slr_worker_train_steps(self.p, n, _d(loss))
        return loss[:n] if ok else None

    def model(self):
        return _model_of(self.handle, self.D)

    def close(self):
        if self.p:
            self.L.slr_worker_destroy(self.p)
            self.p = _p()
