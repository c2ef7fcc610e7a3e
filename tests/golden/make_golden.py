"""Generate tests/golden/lr_golden.npz — fixtures for the BUILD-DEFINED minibatch step.

There is no reference implementation of a gradient step to import (the reference's
training is `model[i] += 1`, src/worker.cc:225-227), so these vectors come from an
independent numpy fp64 restatement of SURVEY.md Appendix D, not from the C oracle:
they pin the C oracle (and through it the CUDA path) against a second implementation.

    python tests/golden/make_golden.py        # rewrites lr_golden.npz
"""
import os

import numpy as np

MASK = (1 << 64) - 1


def splitmix64(x):
    z = (x + 0x9E3779B97F4A7C15) & MASK
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & MASK
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & MASK
    return z ^ (z >> 31)


def synth(row0, nrows, F, seed=0x5EED):
    X = np.empty((nrows, F), dtype=np.float32)
    for r in range(nrows):
        for c in range(F):
            u = splitmix64(seed ^ (((row0 + r) * F + c) & MASK))
            X[r, c] = np.float32(u >> 40) * np.float32(1.0 / 16777216.0) * np.float32(2) - np.float32(1)
    y = (X[:, 0] > 0).astype(np.float32)
    return X, y


def lr_steps(X, y, B, t0, n_steps, w, eta):
    n = X.shape[0]
    Xd, yd = X.astype(np.float64), y.astype(np.float64)
    losses = []
    for s in range(n_steps):
        rows = ((t0 + s) * B + np.arange(B)) % n
        xb, yb = Xd[rows], yd[rows]
        z = xb @ w
        p = 1.0 / (1.0 + np.exp(-z))
        losses.append(np.mean(np.maximum(z, 0) + np.log1p(np.exp(-np.abs(z))) - yb * z))
        w = w - eta * (xb.T @ (p - yb)) / B
    return w, np.array(losses)


CASES = {  # name: (n_rows, F, B, n_steps, eta)
    "tiny": (64, 8, 16, 6, 0.1),
    "ragged": (50, 12, 7, 20, 0.5),       # B does not divide n_rows: minibatches wrap
    "one_row_batches": (5, 4, 1, 11, 1.0),
    "batch_gt_shard": (6, 16, 16, 3, 0.1),  # B > n_rows: rows repeat inside a batch
    "f128": (96, 128, 32, 8, 0.1),
}

if __name__ == "__main__":
    out = {}
    X, _ = synth(0, 4, 8)
    out["synth_rows0_4_F8"] = X
    X, _ = synth(1 << 20, 2, 1024)
    out["synth_row2p20_F1024_first8"] = X[:, :8].copy()
    for name, (n, F, B, steps, eta) in CASES.items():
        X, y = synth(0, n, F)
        w, losses = lr_steps(X, y, B, 0, steps, np.zeros(F), eta)
        out[f"{name}_w"] = w
        out[f"{name}_loss"] = losses
    np.savez(os.path.join(os.path.dirname(os.path.abspath(__file__)), "lr_golden.npz"), **out)
    print({k: v.shape for k, v in out.items()})
