"""CPU-only: the C-ABI libraries load and export every symbol include/*.h declares; host
logic that needs no GPU (wire codec, file server, master membership) against the oracle.
No compute call is made here — on a box without a GPU sl_create must fail loudly."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared(header):
    txt = open(os.path.join(ROOT, "include", header)).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(slr?_[a-z0-9_]+)\s*\(", txt)))


@pytest.fixture(scope="module")
def libs():
    from serverless_learn_b200 import build
    build.build()
    from serverless_learn_b200 import _lib, roles
    return _lib.load(), roles.load()


def test_gpu_abi_exports_every_declared_symbol(libs):
    names = _declared("sl_gpu.h")
    assert len(names) >= 25 and "sl_lr_step_noalloc" in names and "sl_merge" in names
    for n in names:
        assert hasattr(libs[0], n), n


def test_roles_abi_exports_every_declared_symbol(libs):
    names = [n for n in _declared("sl_roles.h") if n.startswith("slr_")]
    assert len(names) >= 30
    for n in names:
        assert hasattr(libs[1], n), n


def test_python_binding_covers_the_headers(libs):
    from serverless_learn_b200 import _lib, roles
    assert set(_declared("sl_gpu.h")) == set(_lib._SIGS)
    assert {n for n in _declared("sl_roles.h") if n.startswith("slr_")} == set(roles._SIGS)


def test_no_cpu_fallback(libs):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import serverless_learn_b200 as sl
    with pytest.raises(sl.SlError) as e:
        sl.Replica(8, 4)
    assert "no CPU fallback" in str(e.value)
    from serverless_learn_b200 import roles
    with pytest.raises(RuntimeError):
        roles.Worker("w0", 8, 4)


def test_product_never_imports_the_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "serverless_learn_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cc", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                for needle in ("import oracle", "from oracle", "sl_oracle", "oracle/", "libsl_ref", "orc_"):
                    assert needle not in src, (f, needle)


# ---- wire codec (src/protos/serverless_learn.proto:81-83) -------------------------------
def test_wire_codec_matches_oracle(libs, orc):
    L = libs[1]
    dp = C.POINTER(C.c_double)
    rng = np.random.default_rng(0)
    for n in (0, 1, 3, 16, 1024, 16384):
        d = rng.standard_normal(n)
        buf = C.create_string_buffer(8 * n + 16)
        w = L.slr_update_encode(d.ctypes.data_as(dp), n, buf, len(buf))
        assert buf.raw[:w] == orc.update_encode(d)
        out = np.empty(max(n, 1))
        assert L.slr_update_decode(buf, w, out.ctypes.data_as(dp), out.size) == n
        assert np.array_equal(out[:n], d)
    d = np.array([1.0, -2.5, 0.0])
    buf = C.create_string_buffer(64)
    w = L.slr_update_encode(d.ctypes.data_as(dp), 3, buf, 64)
    assert buf.raw[:w].hex() == "0a18000000000000f03f00000000000004c00000000000000000"   # B.5
    assert L.slr_update_encode(d.ctypes.data_as(dp), 3, buf, 10) == -1
    assert L.slr_update_decode(b"\x0a\x07" + b"\0" * 7, 9, d.ctypes.data_as(dp), 3) == -1
    # malformed lengths from the wire: 2^63 (negative once narrowed), 2^32 + 8 (truncates to 1
    # element), and a length that only exceeds the output capacity — all refused, nothing copied
    guard = np.full(4, 7.0)
    for varint in (b"\x80" * 9 + b"\x01", b"\x88\x80\x80\x80\x10", b"\x20"):
        msg = b"\x0a" + varint + b"\0" * 64
        assert L.slr_update_decode(msg, len(msg), guard.ctypes.data_as(dp), 3) == -1
        assert np.array_equal(guard, np.full(4, 7.0))


# ---- file server (src/file_server.cc) ----------------------------------------------------
def test_file_server_file0_and_chunks(libs, orc, file0):
    from serverless_learn_b200 import roles
    fs = roles.FileServer()
    assert fs.file_length() == 100_000_000
    want = file0[0]
    for k in (0, 1, 37, 99):
        ch = fs.chunk(k)
        assert ch.size == 1_000_000
        assert np.array_equal(ch, want[k * 1_000_000:(k + 1) * 1_000_000])
    assert fs.chunk(100).size == 0 and fs.chunk(-1).size == 0
    whole = np.concatenate([fs.chunk(k) for k in range(100)])
    assert orc.fnv1a64(whole) == 0xFD6032AAAB1079CF
    d = roles.Directory()
    assert fs.do_push(d, "nobody:1", 0) == 0        # unreachable recipient: ok=false
    assert fs.do_push(d, "nobody:1", 1) == -1       # file_num != 0 (reference: exit(1))
    fs.close()
    d.close()


def test_file_server_synthetic_shards(libs, orc):
    from serverless_learn_b200 import roles
    fs = roles.FileServer()
    assert not fs.add_synthetic_shard(0, 0, 10, 8)          # file 0 stays the reference's dummy file
    assert fs.add_synthetic_shard(3, 100, 1000, 260)
    assert fs.file_length_of(3) == 1000 * 261 * 4 and fs.file_length_of(2) == -1
    stream = np.concatenate([fs.chunk_of(3, k) for k in range(2)])
    assert fs.chunk_of(3, 2).size == 0 and stream.size == 1000 * 261 * 4
    X, y = orc.synth(100, 1000, 260)
    raw = stream.view(np.float32)
    assert np.array_equal(raw[:260000].reshape(1000, 260), X) and np.array_equal(raw[260000:], y)
    d = roles.Directory()
    assert fs.do_push(d, "nobody:1", 2) == -1 and fs.do_push(d, "nobody:1", 3) == 0
    fs.close(); d.close()


def test_file_server_described_shard_header(libs, orc):
    # §8f n2: the 64-byte sl_shard_header (include/sl_gpu.h) in front of the payload
    import struct
    from serverless_learn_b200 import roles
    fs = roles.FileServer()
    assert fs.add_described_shard(5, 777, 300, 64)
    n = fs.file_length_of(5)
    assert n == 64 + 300 * 65 * 4
    stream = np.concatenate([fs.chunk_of(5, k) for k in range(1)])
    assert stream.size == n
    magic, version, F, n_rows, row0, flags, _r0, bsum, wsum, _r1 = struct.unpack("<8sIIQQIIQQQ", stream[:64].tobytes())
    assert (magic, version, F, n_rows, row0, flags) == (b"SLSHARD\x01", 1, 64, 300, 777, 1)
    payload = stream[64:]
    assert bsum == orc.bytesum(payload) and wsum == orc.weighted_sum(payload)
    X, y = orc.synth(777, 300, 64)
    raw = payload.view(np.float32)
    assert np.array_equal(raw[:300 * 64].reshape(300, 64), X) and np.array_equal(raw[300 * 64:], y)
    fs.close()


def test_cpp_job_program_fails_loudly_without_a_gpu(libs):
    # examples/sl_job.cc is built by build(); with no CUDA device it must stop with the library's message
    import subprocess
    exe = os.path.join(ROOT, "serverless_learn_b200", "sl_job")
    assert os.path.exists(exe), "build() did not produce sl_job"
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("a GPU is present")
    except ImportError:
        pass
    out = subprocess.run([exe, "--workers", "2", "--rows", "64", "--features", "16", "--batch", "8"],
                         capture_output=True, text=True, timeout=120)
    assert out.returncode == 1 and "no CUDA device" in out.stderr, (out.returncode, out.stderr)
