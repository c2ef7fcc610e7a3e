"""CPU oracle for the serverless_learn hot path — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this package, and only as the checker or the timed CPU baseline.
The product package (serverless_learn_b200) never imports it.

`Oracle` wraps oracle/libsl_oracle.so (plain C restatement, see sl_oracle.h for the
reference file:line each function follows).  `Ref` wraps oracle/_ref/libsl_ref.so,
the reference's own hot-path lines compiled by oracle/build_ref.sh; it exists only
where that library was built (a container with /root/reference) or travelled.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
FILE0_LENGTH = 100_000_000  # src/file_server.cc:40
CHUNK_SIZE = 1_000_000      # src/file_server.cc:46
SEED = 0x5EED

_dp = C.POINTER(C.c_double)
_fp = C.POINTER(C.c_float)
_u8p = C.POINTER(C.c_uint8)


def build(force: bool = False) -> None:
    """Compile the oracle (and oracle/_ref when /root/reference is present)."""
    so = os.path.join(_HERE, "libsl_oracle.so")
    srcs = [os.path.join(_HERE, f) for f in ("sl_oracle.c", "sl_oracle_lr.c", "sl_oracle.h")]
    stale = force or not os.path.exists(so) or any(
        os.path.getmtime(s) > os.path.getmtime(so) for s in srcs)
    if stale:
        subprocess.run(["make", "-C", _HERE, "libsl_oracle.so"], check=True,
                       stdout=subprocess.DEVNULL)
    ref = os.path.join(_HERE, "_ref", "libsl_ref.so")
    if (force or not os.path.exists(ref)) and os.path.isdir(
            os.environ.get("SL_REFERENCE_DIR", "/root/reference")):
        subprocess.run(["bash", os.path.join(_HERE, "build_ref.sh")], check=True,
                       stdout=subprocess.DEVNULL)


def _d(a):
    return a.ctypes.data_as(_dp)


def _f(a):
    return a.ctypes.data_as(_fp)


class Oracle:
    def __init__(self):
        build()
        L = C.CDLL(os.path.join(_HERE, "libsl_oracle.so"))
        L.orc_file0.restype = C.c_int
        L.orc_file0.argtypes = [_u8p, C.c_size_t, C.POINTER(C.c_int64), C.c_int]
        for n in ("orc_fnv1a64", "orc_bytesum"):
            getattr(L, n).restype = C.c_uint64
            getattr(L, n).argtypes = [_u8p, C.c_size_t]
        L.orc_weighted_sum.restype = C.c_uint64
        L.orc_weighted_sum.argtypes = [_u8p, C.c_size_t, C.c_uint64]
        L.orc_chunk_count.restype = C.c_int64
        L.orc_chunk_count.argtypes = [C.c_int64, C.c_int64]
        L.orc_chunk_len.restype = C.c_int64
        L.orc_chunk_len.argtypes = [C.c_int64, C.c_int64, C.c_int64]
        L.orc_merge_callee.restype = C.c_int
        L.orc_merge_callee.argtypes = [_dp, _dp, C.c_int, _dp, C.c_int, C.c_double, _dp]
        L.orc_delta.restype = None
        L.orc_delta.argtypes = [_dp, _dp, C.c_int, _dp]
        L.orc_apply_reply.restype = C.c_int
        L.orc_apply_reply.argtypes = [_dp, _dp, C.c_int, _dp, C.c_int, C.c_double]
        L.orc_train_tick.restype = None
        L.orc_train_tick.argtypes = [_dp, C.c_int]
        L.orc_update_encode.restype = C.c_long
        L.orc_update_encode.argtypes = [_dp, C.c_int, _u8p, C.c_long]
        L.orc_update_decode.restype = C.c_int
        L.orc_update_decode.argtypes = [_u8p, C.c_long, _dp, C.c_int]
        L.orc_splitmix64.restype = C.c_uint64
        L.orc_splitmix64.argtypes = [C.c_uint64]
        L.orc_synth_fill.restype = None
        L.orc_synth_fill.argtypes = [_fp, _fp, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64]
        L.orc_lr_grad.restype = C.c_double
        L.orc_lr_grad.argtypes = [_fp, _fp, C.c_int64, C.c_int, C.c_int, C.c_int64, _dp, _dp, _dp]
        L.orc_lr_steps.restype = None
        L.orc_lr_steps.argtypes = [_fp, _fp, C.c_int64, C.c_int, C.c_int, C.c_int64, C.c_int, _dp,
                                   C.c_double, _dp]
        L.orc_lr_steps_sync.restype = None
        L.orc_lr_steps_sync.argtypes = [C.POINTER(_fp), C.POINTER(_fp), C.POINTER(C.c_int64),
                                        C.c_int, C.c_int, C.c_int, C.c_int64, C.c_int, _dp,
                                        C.c_double, _dp, C.c_int]
        L.orc_lr_bench_workers.restype = C.c_double
        L.orc_lr_bench_workers.argtypes = [C.POINTER(_fp), C.POINTER(_fp), C.POINTER(C.c_int64),
                                           C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(_dp),
                                           C.c_double]
        L.orc_max_threads.restype = C.c_int
        self.L = L

    # ---- file 0 -----------------------------------------------------------
    def file0(self, n: int = FILE0_LENGTH, want_rejections: bool = False):
        out = np.empty(n, dtype=np.uint8)
        rej = np.zeros(64, dtype=np.int64)
        k = self.L.orc_file0(out.ctypes.data_as(_u8p), n, rej.ctypes.data_as(C.POINTER(C.c_int64)), 64)
        return (out, rej[:k].copy()) if want_rejections else out

    def fnv1a64(self, b: np.ndarray) -> int:
        b = np.ascontiguousarray(b, dtype=np.uint8)
        return int(self.L.orc_fnv1a64(b.ctypes.data_as(_u8p), b.size))

    def bytesum(self, b: np.ndarray) -> int:
        b = np.ascontiguousarray(b, dtype=np.uint8)
        return int(self.L.orc_bytesum(b.ctypes.data_as(_u8p), b.size))

    def weighted_sum(self, b: np.ndarray, base: int = 0) -> int:
        b = np.ascontiguousarray(b, dtype=np.uint8)
        return int(self.L.orc_weighted_sum(b.ctypes.data_as(_u8p), b.size, base))

    def chunk_lens(self, length: int, chunk: int = CHUNK_SIZE):
        n = self.L.orc_chunk_count(length, chunk)
        return [int(self.L.orc_chunk_len(length, chunk, k)) for k in range(n)]

    # ---- exchange ---------------------------------------------------------
    def merge_callee(self, model, old, delta_in, lam=0.5):
        """In place on model/old; returns reply (or None if lengths differ)."""
        delta_in = np.ascontiguousarray(delta_in, dtype=np.float64)
        reply = np.empty_like(model)
        rc = self.L.orc_merge_callee(_d(model), _d(old), model.size, _d(delta_in), delta_in.size,
                                     lam, _d(reply))
        return reply if rc == 0 else None

    def delta(self, model, old):
        out = np.empty_like(model)
        self.L.orc_delta(_d(model), _d(old), model.size, _d(out))
        return out

    def apply_reply(self, model, old, back, lam=0.5):
        back = np.ascontiguousarray(back, dtype=np.float64)
        return self.L.orc_apply_reply(_d(model), _d(old), model.size, _d(back), back.size, lam)

    def train_tick(self, model):
        self.L.orc_train_tick(_d(model), model.size)

    # ---- wire -------------------------------------------------------------
    def update_encode(self, delta) -> bytes:
        delta = np.ascontiguousarray(delta, dtype=np.float64)
        buf = np.empty(delta.size * 8 + 16, dtype=np.uint8)
        n = self.L.orc_update_encode(_d(delta), delta.size, buf.ctypes.data_as(_u8p), buf.size)
        assert n >= 0
        return buf[:n].tobytes()

    def update_decode(self, raw: bytes):
        buf = np.frombuffer(raw, dtype=np.uint8).copy() if raw else np.zeros(1, np.uint8)
        out = np.empty(max(1, len(raw) // 8 + 1), dtype=np.float64)
        n = self.L.orc_update_decode(buf.ctypes.data_as(_u8p), len(raw), _d(out), out.size)
        return None if n < 0 else out[:n].copy()

    # ---- synthetic data ----------------------------------------------------
    def synth(self, row0: int, nrows: int, F: int, seed: int = SEED):
        X = np.empty((nrows, F), dtype=np.float32)
        y = np.empty(nrows, dtype=np.float32)
        self.L.orc_synth_fill(_f(X), _f(y), row0, nrows, F, seed)
        return X, y

    # ---- logistic regression (build-defined) --------------------------------
    def lr_grad(self, X, y, B, t, w):
        g = np.empty(X.shape[1], dtype=np.float64)
        loss = self.L.orc_lr_grad(_f(X), _f(y), X.shape[0], X.shape[1], B, t, _d(w), _d(g), None)
        return g, float(loss)

    def lr_steps(self, X, y, B, t0, n_steps, w, eta):
        """Updates w in place; returns per-step losses."""
        loss = np.empty(n_steps, dtype=np.float64)
        self.L.orc_lr_steps(_f(X), _f(y), X.shape[0], X.shape[1], B, t0, n_steps, _d(w), eta,
                            _d(loss))
        return loss

    def _ptr_arrays(self, Xs, ys):
        W = len(Xs)
        xp = (_fp * W)(*[_f(x) for x in Xs])
        yp = (_fp * W)(*[_f(y) for y in ys])
        nr = (C.c_int64 * W)(*[x.shape[0] for x in Xs])
        return xp, yp, nr

    def lr_steps_sync(self, Xs, ys, B, t0, n_steps, w, eta, threads=1):
        xp, yp, nr = self._ptr_arrays(Xs, ys)
        loss = np.empty(n_steps, dtype=np.float64)
        self.L.orc_lr_steps_sync(xp, yp, nr, len(Xs), Xs[0].shape[1], B, t0, n_steps, _d(w), eta,
                                 _d(loss), threads)
        return loss

    def lr_bench_workers(self, Xs, ys, B, n_steps, ws, eta) -> float:
        xp, yp, nr = self._ptr_arrays(Xs, ys)
        wp = (_dp * len(ws))(*[_d(w) for w in ws])
        return float(self.L.orc_lr_bench_workers(xp, yp, nr, len(Xs), Xs[0].shape[1], B, n_steps,
                                                 wp, eta))

    def max_threads(self) -> int:
        return int(self.L.orc_max_threads())


class Ref:
    """The reference's own lines (oracle/_ref/libsl_ref.so).  `Ref.available()` is
    False where the library was neither built nor shipped."""

    PATH = os.path.join(_HERE, "_ref", "libsl_ref.so")

    @classmethod
    def available(cls) -> bool:
        return os.path.exists(cls.PATH)

    def __init__(self):
        L = C.CDLL(self.PATH)
        L.ref_file0_length.restype = C.c_long
        L.ref_chunk_size.restype = C.c_long
        L.ref_learn_rate.restype = C.c_double
        L.ref_file0.restype = C.c_long
        L.ref_file0.argtypes = [_u8p, C.c_long]
        L.ref_merge.restype = C.c_int
        L.ref_merge.argtypes = [_dp, _dp, C.c_int, _dp, C.c_int, _dp, C.c_int]
        L.ref_delta.restype = C.c_int
        L.ref_delta.argtypes = [_dp, _dp, C.c_int, _dp]
        L.ref_apply.restype = C.c_int
        L.ref_apply.argtypes = [_dp, _dp, C.c_int, _dp, C.c_int]
        L.ref_tick.restype = None
        L.ref_tick.argtypes = [_dp, C.c_int]
        self.L = L

    def file0(self):
        n = self.L.ref_file0_length()
        out = np.empty(n, dtype=np.uint8)
        got = self.L.ref_file0(out.ctypes.data_as(_u8p), n)
        assert got == n
        return out

    def merge(self, model, old, delta_in):
        delta_in = np.ascontiguousarray(delta_in, dtype=np.float64)
        reply = np.empty_like(model)
        n = self.L.ref_merge(_d(model), _d(old), model.size, _d(delta_in), delta_in.size,
                             _d(reply), reply.size)
        return reply, n

    def delta(self, model, old):
        out = np.empty_like(model)
        n = self.L.ref_delta(_d(model), _d(old), model.size, _d(out))
        return out[:n]

    def apply(self, model, old, back):
        back = np.ascontiguousarray(back, dtype=np.float64)
        return self.L.ref_apply(_d(model), _d(old), model.size, _d(back), back.size)

    def tick(self, model):
        self.L.ref_tick(_d(model), model.size)
