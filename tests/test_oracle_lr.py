"""BUILD-DEFINED logistic-regression oracle (oracle/sl_oracle_lr.c) vs the committed
golden vectors (independent numpy restatement, tests/golden/make_golden.py) and vs
size-independent properties.  Parity here is unpinned by the reference: it has no
gradient step (src/worker.cc:221-231 adds 1)."""
import importlib.util
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = np.load(os.path.join(HERE, "golden", "lr_golden.npz"))
_spec = importlib.util.spec_from_file_location("make_golden", os.path.join(HERE, "golden", "make_golden.py"))
mg = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(mg)


def test_synth_matches_golden(orc):
    X, y = orc.synth(0, 4, 8)
    assert np.array_equal(X, GOLD["synth_rows0_4_F8"])
    assert np.array_equal(y, (X[:, 0] > 0).astype(np.float32))
    X, _ = orc.synth(1 << 20, 2, 1024)
    assert np.array_equal(X[:, :8], GOLD["synth_row2p20_F1024_first8"])
    assert X.min() >= -1.0 and X.max() < 1.0


def test_synth_is_counter_based(orc):
    full, yf = orc.synth(0, 40, 24)
    part, yp = orc.synth(13, 9, 24)
    assert np.array_equal(full[13:22], part) and np.array_equal(yf[13:22], yp)


@pytest.mark.parametrize("name", list(mg.CASES))
def test_lr_steps_match_golden(orc, name):
    n, F, B, steps, eta = mg.CASES[name]
    X, y = orc.synth(0, n, F)
    w = np.zeros(F)
    loss = orc.lr_steps(X, y, B, 0, steps, w, eta)
    np.testing.assert_allclose(w, GOLD[f"{name}_w"], rtol=0, atol=1e-13)
    np.testing.assert_allclose(loss, GOLD[f"{name}_loss"], rtol=1e-13)


def test_first_loss_is_ln2(orc):
    X, y = orc.synth(0, 256, 32)
    w = np.zeros(32)
    loss = orc.lr_steps(X, y, 64, 0, 1, w, 0.1)
    assert abs(loss[0] - np.log(2.0)) < 1e-15


def test_steps_compose(orc):
    X, y = orc.synth(0, 100, 16)
    w1 = np.zeros(16)
    l1 = orc.lr_steps(X, y, 8, 0, 30, w1, 0.3)
    w2 = np.zeros(16)
    la = orc.lr_steps(X, y, 8, 0, 11, w2, 0.3)
    lb = orc.lr_steps(X, y, 8, 11, 19, w2, 0.3)
    assert np.array_equal(w1, w2) and np.array_equal(l1, np.concatenate([la, lb]))


def test_loss_decreases_on_learnable_labels(orc):
    X, y = orc.synth(0, 4096, 64)
    w = np.zeros(64)
    loss = orc.lr_steps(X, y, 256, 0, 200, w, 0.5)
    assert loss[-1] < 0.5 * loss[0]
    assert w[0] > 5 * np.abs(w[1:]).max()      # y depends on column 0 only


def test_sync_workers_equal_one_big_batch(orc):
    # W workers averaging each step == one worker with the W minibatches concatenated
    W, n, F, B, steps = 4, 64, 12, 8, 10
    Xs, ys = zip(*[orc.synth(k * n, n, F) for k in range(W)])
    w = np.zeros(F)
    loss = orc.lr_steps_sync(list(Xs), list(ys), B, 0, steps, w, 0.2, threads=1)
    wt = np.zeros(F)
    lt = orc.lr_steps_sync(list(Xs), list(ys), B, 0, steps, wt, 0.2, threads=4)
    assert np.array_equal(w, wt) and np.array_equal(loss, lt)
    # numpy: interleave so that global step s uses rows [s*B,(s+1)*B) of every shard
    wn = np.zeros(F)
    for s in range(steps):
        g = np.zeros(F)
        for k in range(W):
            gk, _ = orc.lr_grad(Xs[k], ys[k], B, s, wn)
            g += gk
        wn = wn - 0.2 * g / W
    np.testing.assert_allclose(w, wn, atol=1e-14)


def test_bench_workers_runs(orc):
    Xs, ys = zip(*[orc.synth(k * 32, 32, 8) for k in range(2)])
    ws = [np.zeros(8), np.zeros(8)]
    dt = orc.lr_bench_workers(list(Xs), list(ys), 4, 5, ws, 0.1)
    assert dt >= 0
    ref = np.zeros(8)
    orc.lr_steps(Xs[1], ys[1], 4, 0, 5, ref, 0.1)
    assert np.array_equal(ws[1], ref)
