"""Allocation-free step (sl_lr_step_noalloc) against the regular entry point.

Collected last on purpose (file name): a run that stops at the first failure has then
already run every other test.  The entry point is for callers that may not allocate device
memory — worker replicas sharing one GPU, where cudaMalloc would stall the neighbours, and
workers whose shard fills HBM — so it works out of a fixed scratch array inside the library.
Checked here: two calls in a row (the second reuses the scratch the first left behind) give
the same weights and losses as two steps of sl_train_steps' simple path, within the
tolerance of the other GPU tests, and oversize shapes are refused.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

W_TOL = 1e-5
L_TOL = 1e-5


def test_noalloc_step_matches_train_steps(orc):
    import torch

    import serverless_learn_b200 as sl
    from serverless_learn_b200 import _lib

    L = _lib.load()
    n, F, B, eta = 700, 260, 100, 0.2
    X, y = orc.synth(0, n, F)

    with sl.Replica(F, B, eta=eta) as r:
        r.set_option("path", 2)
        r.load_shard(X, y)
        want_loss = r.train_steps(2)
        want_w = r.model()

    dev = torch.device("cuda", 0)
    dX = torch.from_numpy(X).to(dev)
    dy = torch.from_numpy(y).to(dev)
    dw = torch.zeros(F, dtype=torch.float64, device=dev)
    dl = torch.zeros(2, dtype=torch.float64, device=dev)
    torch.cuda.synchronize()
    for t in range(2):      # second call runs on the scratch contents of the first
        rc = L.sl_lr_step_noalloc(0, dX.data_ptr(), dy.data_ptr(), n, F, B, t, eta, dw.data_ptr(),
                                  dl.data_ptr() + 8 * t, None)
        assert rc == 0, L.sl_last_error(None)
    torch.cuda.synchronize()
    got_w = dw.cpu().numpy()
    got_loss = dl.cpu().numpy()
    assert np.max(np.abs(got_w - want_w)) / np.max(np.abs(want_w)) <= W_TOL
    assert np.max(np.abs(got_loss - want_loss) / np.abs(want_loss)) <= L_TOL
    # and against the oracle itself
    w = np.zeros(F)
    lo = orc.lr_steps(X, y, B, 0, 2, w, eta)
    assert np.max(np.abs(got_w - w)) / np.max(np.abs(w)) <= W_TOL
    assert np.max(np.abs(got_loss - lo) / lo) <= L_TOL


def test_noalloc_rejects_oversize():
    import torch

    from serverless_learn_b200 import _lib

    L = _lib.load()
    buf = torch.zeros(16, dtype=torch.float64, device="cuda:0")
    p = buf.data_ptr()
    assert L.sl_lr_step_noalloc(0, p, p, 10, 4, 8193, 0, 0.1, p, None, None) == -1
    assert b"8192" in L.sl_last_error(None)
    assert L.sl_lr_step_noalloc(0, p, p, 10, 4100, 8, 0, 0.1, p, None, None) == -1
    assert L.sl_lr_step_noalloc(0, None, p, 10, 4, 8, 0, 0.1, p, None, None) == -1


def test_noalloc_rejects_misaligned_rows():
    """With F a multiple of 4 the rows are read 16 bytes at a time: a view 4 bytes into a buffer is
    refused up front instead of faulting on the device; F not a multiple of 4 takes any alignment."""
    import torch

    from serverless_learn_b200 import _lib

    L = _lib.load()
    buf = torch.zeros(1024, dtype=torch.float32, device="cuda:0")
    w = torch.zeros(8, dtype=torch.float64, device="cuda:0")
    assert L.sl_lr_step_noalloc(0, buf.data_ptr() + 4, buf.data_ptr(), 10, 8, 4, 0, 0.1, w.data_ptr(), None, None) == -1
    assert b"aligned" in L.sl_last_error(None)
    assert L.sl_lr_step_noalloc(0, buf.data_ptr() + 4, buf.data_ptr(), 10, 7, 4, 0, 0.1, w.data_ptr(), None, None) == 0
    torch.cuda.synchronize()


def test_noalloc_at_its_limits(orc):
    """B = 8192 and F = 4096, the largest shape the built-in scratch takes, against the oracle."""
    import torch

    from serverless_learn_b200 import _lib

    L = _lib.load()
    n, F, B, eta = 8192, 4096, 8192, 0.1
    X, y = orc.synth(0, n, F)
    dev = torch.device("cuda", 0)
    dX, dy = torch.from_numpy(X).to(dev), torch.from_numpy(y).to(dev)
    dw = torch.zeros(F, dtype=torch.float64, device=dev)
    dl = torch.zeros(2, dtype=torch.float64, device=dev)
    for t in range(2):
        assert L.sl_lr_step_noalloc(0, dX.data_ptr(), dy.data_ptr(), n, F, B, t, eta, dw.data_ptr(),
                                    dl.data_ptr() + 8 * t, None) == 0, L.sl_last_error(None)
    torch.cuda.synchronize()
    w = np.zeros(F)
    lo = orc.lr_steps(X, y, B, 0, 2, w, eta)
    assert np.max(np.abs(dw.cpu().numpy() - w)) / np.max(np.abs(w)) <= W_TOL
    assert np.max(np.abs(dl.cpu().numpy() - lo) / lo) <= L_TOL
