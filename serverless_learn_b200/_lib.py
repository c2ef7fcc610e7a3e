"""ctypes binding of libsl_gpu.so (include/sl_gpu.h).  No fallback: if the library is
missing or the GPU is absent the caller gets an exception, never a CPU path."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SL_GPU_LIB") or os.path.join(HERE, "libsl_gpu.so")

SL_OK, SL_EINVAL, SL_ECUDA, SL_ESTATE, SL_ENOMEM, SL_ETIMEOUT, SL_ENCCL = 0, -1, -2, -3, -4, -5, -6

_dp = C.POINTER(C.c_double)
_h = C.c_void_p

_SIGS = {
    "sl_create": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.POINTER(_h)]),
    "sl_destroy": (C.c_int, [_h]),
    "sl_last_error": (C.c_char_p, [_h]),
    "sl_set_option": (C.c_int, [_h, C.c_char_p, C.c_longlong]),
    "sl_get_info": (C.c_longlong, [_h, C.c_char_p]),
    "sl_debug_read": (C.c_int, [_h, C.POINTER(C.c_ulonglong), C.c_size_t]),
    "sl_shard_begin": (C.c_int, [_h, C.c_size_t]),
    "sl_shard_append": (C.c_int, [_h, C.c_void_p, C.c_size_t]),
    "sl_shard_end": (C.c_int, [_h, C.POINTER(C.c_longlong)]),
    "sl_shard_checksum": (C.c_int, [_h, C.POINTER(C.c_ulonglong), C.POINTER(C.c_ulonglong)]),
    "sl_shard_read": (C.c_int, [_h, C.c_size_t, C.c_void_p, C.c_size_t]),
    "sl_shard_bytes": (C.c_longlong, [_h]),
    "sl_shard_synth": (C.c_int, [_h, C.c_longlong, C.c_ulonglong, C.c_ulonglong]),
    "sl_train_tick": (C.c_int, [_h]),
    "sl_train_steps": (C.c_int, [_h, C.c_int, _dp]),
    "sl_last_train_ms": (C.c_float, [_h]),
    "sl_merge": (C.c_int, [_h, _dp, C.c_int, _dp, C.c_int]),
    "sl_delta_snapshot": (C.c_int, [_h, _dp, C.c_int]),
    "sl_apply_reply": (C.c_int, [_h, _dp, C.c_int]),
    "sl_get_model": (C.c_int, [_h, _dp, C.c_int]),
    "sl_get_old": (C.c_int, [_h, _dp, C.c_int]),
    "sl_set_state": (C.c_int, [_h, _dp, _dp, C.c_int]),
    "sl_gossip_p2p": (C.c_int, [_h, _h]),
    "sl_state_export": (C.c_int, [_h, C.c_void_p]),
    "sl_state_attach": (C.c_int, [_h, C.c_void_p, C.c_int, C.c_double, C.POINTER(C.c_int)]),
    "sl_state_detach": (C.c_int, [_h, C.c_int]),
    "sl_state_lend": (C.c_int, [_h]),
    "sl_state_reclaim": (C.c_int, [_h]),
    "sl_gossip_p2p_remote": (C.c_int, [_h, C.c_int]),
    "sl_peer_export": (C.c_int, [_h, C.c_void_p]),
    "sl_peer_attach": (C.c_int, [_h, C.c_int, C.c_int, C.c_void_p]),
    "sl_peer_detach": (C.c_int, [_h]),
    "sl_peer_attach_symmetric": (C.c_int, [_h, C.c_int, C.c_int, C.POINTER(C.c_void_p), C.c_void_p, C.c_size_t]),
    "sl_comm_unique_id": (C.c_int, [C.c_void_p]),
    "sl_comm_init_rank": (C.c_int, [_h, C.c_int, C.c_int, C.c_void_p]),
    "sl_allreduce_avg": (C.c_int, [_h]),
    "sl_train_steps_sync": (C.c_int, [_h, C.c_int, _dp]),
    "sl_train_steps_group": (C.c_int, [C.POINTER(_h), C.c_int, C.c_int, _dp]),
    "sl_train_steps_multi": (C.c_int, [C.POINTER(_h), C.c_int, C.c_int, _dp]),
    "sl_lr_step_noalloc": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_longlong, C.c_int, C.c_int,
                                     C.c_longlong, C.c_double, C.c_void_p, C.c_void_p, C.c_void_p]),
}

_lib = None


def load() -> C.CDLL:
    """Load libsl_gpu.so (built by serverless_learn_b200.build).  Raises if absent."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: run `python -m serverless_learn_b200.build`. "
                "There is no CPU fallback.")
        if "SL_NCCL_LIB" not in os.environ:
            # if PyTorch is installed, make the library pick the NCCL that torch bundles: a later
            # `import torch` in the same process would otherwise meet a second libnccl.so.2
            import importlib.util
            spec = importlib.util.find_spec("nvidia")
            for base in (list(spec.submodule_search_locations) if spec and spec.submodule_search_locations else []):
                cand = os.path.join(base, "nccl", "lib", "libnccl.so.2")
                if os.path.exists(cand):
                    os.environ["SL_NCCL_LIB"] = cand
                    break
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGS.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib
