"""K1c (lr_cluster_kernel: thread-block clusters, distributed-shared-memory reduce-scatter, one
add-with-count hop through L2) against the CPU oracle, through the C ABI.  Same tolerances as
tests/test_gpu_parity.py; the kernel takes the slot of simulate_training (src/worker.cc:221-231).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

W_TOL = 1e-5
L_TOL = 1e-5


@pytest.fixture(scope="module")
def sl():
    import serverless_learn_b200 as m
    return m


def w_err(a, b):
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-12))


def _run(sl, orc, n, F, B, steps, eta, clv=0, clusters=0, split=None, want_loss=True):
    X, y = orc.synth(0, n, F)
    w = np.zeros(F)
    lo = orc.lr_steps(X, y, B, 0, steps, w, eta)
    with sl.Replica(F, B, eta=eta) as r:
        r.set_option("path", 6)
        r.set_option("cl_variant", clv)
        r.set_option("clusters", clusters)
        r.load_shard(X, y)
        parts = [split, steps - split] if split else [steps]
        got = [r.train_steps(k, want_loss=want_loss) for k in parts]
        lg = np.concatenate(got) if want_loss else None
        wg = r.model()
        info = (r.info("path"), r.info("variant"), r.info("grid"))
    el = float(np.max(np.abs(lg - lo) / np.abs(lo))) if want_loss else 0.0
    return w_err(wg, w), el, info, wg, lg


SHAPES = [
    (300, 64, 48),        # fewer rows per step than CTAs: most CTAs only take part in the exchange
    (2000, 1024, 256),    # two rows per CTA
    (700, 256, 100),      # ragged
    (37, 512, 200),       # B > n_rows: rows repeat inside a minibatch, every step wraps the shard
    (9000, 1024, 4096),   # config-3 minibatch: 32 rows per CTA
    (5000, 992, 1111),    # F a multiple of 32 below the maximum, B not a multiple of anything
    (40000, 1024, 16384),  # several passes per row group and step
    (3000, 2048, 600),    # rows wider than 1024 floats: the 16-warp row group (variants 17, 18)
    (5000, 1536, 4096),
]


@pytest.mark.parametrize("clv", list(range(1, 19)))
@pytest.mark.parametrize("shape", SHAPES)
def test_cluster_kernel_matches_the_oracle(sl, orc, clv, shape):
    n, F, B = shape
    if (clv == 4 and F > 512) or (clv <= 16 and F > 1024):
        with pytest.raises(sl.SlError):
            _run(sl, orc, 64, F, 16, 1, 0.2, clv=clv)
        return
    ew, el, info, _, _ = _run(sl, orc, n, F, B, 7, 0.2, clv=clv)
    assert info[0] == 6 and info[1] == 100 + clv
    assert ew <= W_TOL and el <= L_TOL, (ew, el, info)


@pytest.mark.parametrize("F,B,want_path", [(2048, 600, 6), (1536, 4096, 6), (2048, 8192, 1)])
def test_wide_rows_choose_their_kernel(sl, orc, F, B, want_path):
    # nothing forced: rows of 1025..2048 floats take the cluster kernel's 16-warp row group up to three passes per
    # step and the older persistent kernel beyond; either way the oracle's numbers
    n, steps, eta = 2 * B + 100, 5, 0.2
    X, y = orc.synth(0, n, F)
    w = np.zeros(F)
    lo = orc.lr_steps(X, y, B, 0, steps, w, eta)
    with sl.Replica(F, B, eta=eta) as r:
        r.load_shard(X, y)
        lg = r.train_steps(steps)
        assert r.info("path") == want_path, (r.info("path"), r.info("variant"))
        if want_path == 6:
            assert r.info("variant") == 117
        assert w_err(r.model(), w) <= W_TOL and float(np.max(np.abs(lg - lo) / lo)) <= L_TOL


@pytest.mark.parametrize("clusters", [1, 2, 5, 0])
def test_cluster_counts(sl, orc, clusters):
    ew, el, info, _, _ = _run(sl, orc, 3000, 512, 700, 9, 0.3, clusters=clusters)
    if clusters:
        assert info[2] == 8 * clusters      # F = 512 runs variant 4: clusters of 8
    assert ew <= W_TOL and el <= L_TOL, (ew, el, info)


def test_cluster_calls_compose_and_are_deterministic(sl, orc):
    # accumulators are never cleared: a second launch starts from the values the first left behind
    a = _run(sl, orc, 900, 256, 64, 21, 0.2)
    b = _run(sl, orc, 900, 256, 64, 21, 0.2, split=7)      # odd split: the parity continues across launches
    c = _run(sl, orc, 900, 256, 64, 21, 0.2)
    assert np.array_equal(a[3], c[3]) and np.array_equal(a[4], c[4])   # run-to-run identical
    assert np.array_equal(a[3], b[3]) and np.array_equal(a[4], b[4])   # split == unsplit
    d = _run(sl, orc, 900, 256, 64, 21, 0.2, want_loss=False)          # the loss path is optional
    assert np.array_equal(a[3], d[3])


def test_cluster_many_short_calls(sl, orc):
    n, F, B, eta = 2000, 1024, 300, 0.1
    X, y = orc.synth(0, n, F)
    w = np.zeros(F)
    lo = orc.lr_steps(X, y, B, 0, 40, w, eta)
    with sl.Replica(F, B, eta=eta) as r:
        r.load_shard(X, y)
        lg = np.concatenate([r.train_steps(k) for k in (1, 1, 2, 3, 1, 5, 8, 1, 1, 13, 4)])
        assert r.info("path") == 6
        assert w_err(r.model(), w) <= W_TOL and np.max(np.abs(lg - lo) / lo) <= L_TOL
        # launch splitting inside one call gives the same bits
        wa = r.model()
    with sl.Replica(F, B, eta=eta) as r:
        r.load_shard(X, y)
        r.set_option("steps_per_launch", 7)
        lg2 = r.train_steps(40)
        assert np.array_equal(r.model(), wa) and np.array_equal(lg2, lg)


def test_cluster_agrees_with_the_persistent_kernel(sl, orc):
    n, F, B, steps = 1 << 15, 1024, 4096, 24
    res = {}
    for path in (6, 1):
        with sl.Replica(F, B) as r:
            r.shard_synth(n)
            r.set_option("path", path)
            res[path] = (r.train_steps(steps), r.model())
            assert r.info("path") == path
    assert w_err(res[6][1], res[1][1]) <= W_TOL
    assert np.max(np.abs(res[6][0] - res[1][0]) / res[1][0]) <= L_TOL


def test_cluster_gradient_out_of_range_is_reported(sl, orc):
    # features scaled by 1e6: a partial gradient leaves the fixed-point range; the call must say so
    n, F, B = 4096, 1024, 4096
    X, y = orc.synth(0, n, F)
    X = (X * np.float32(1e6)).astype(np.float32)
    with sl.Replica(F, B, eta=1e-9) as r:
        r.set_option("path", 6)
        r.load_shard(X, y)
        with pytest.raises(sl.SlError) as ei:
            r.train_steps(2)
        assert ei.value.code == -1 and "range" in str(ei.value)


@pytest.mark.parametrize("W,n,F,B", [(4, 640, 64, 48), (4, 1000, 1024, 256), (3, 500, 256, 100), (8, 96, 32, 7),
                                     (4, 20000, 1024, 4096)])
def test_cluster_group_of_workers(sl, orc, W, n, F, B):
    steps, eta = 9, 0.2
    shards = [orc.synth(k * n, n, F) for k in range(W)]
    w = np.zeros(F)
    lo = orc.lr_steps_sync([s[0] for s in shards], [s[1] for s in shards], B, 0, steps, w, eta)
    reps = [sl.Replica(F, B, eta=eta) for _ in range(W)]
    try:
        for r, (X, y) in zip(reps, shards):
            r.load_shard(X, y)
        reps[0].set_option("path", 6)
        lg = np.concatenate([sl.Replica.train_steps_group(reps, 4), sl.Replica.train_steps_group(reps, steps - 4)])
        assert reps[0].info("path") == 6
        for r in reps:
            assert np.array_equal(r.model(), reps[0].model())
        assert w_err(reps[0].model(), w) <= W_TOL and np.max(np.abs(lg - lo) / lo) <= L_TOL
    finally:
        for r in reps:
            r.close()


def test_cluster_full_size_properties(sl):
    # full config-3 size (2^20 x 1024, 4 GiB resident): properties that need no oracle
    n, F, B = 1 << 20, 1024, 4096
    with sl.Replica(F, B) as r:
        r.shard_synth(n)
        l1 = r.train_steps(300)          # crosses the epoch boundary at step 256
        w1 = r.model()
        assert r.info("path") == 6
        assert np.all(np.isfinite(l1)) and l1[-1] < 0.6 * l1[0]
        assert abs(l1[0] - np.log(2)) < 1e-6
        assert w1[0] > 5 * np.abs(w1[1:]).max()     # labels depend on column 0 only
        r.set_option("step", 0)
        r.set_state(np.zeros(F), np.zeros(F))
        l2 = r.train_steps(300)
        assert np.array_equal(l1, l2) and np.array_equal(w1, r.model())
