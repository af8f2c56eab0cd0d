"""CPU tests: MNIST loader semantics and the numpy oracle."""
import numpy as np
import pytest

from distributedtensorflowexample_b200.data import DataSet, dense_to_one_hot, read_data_sets, synthetic_arrays
from distributedtensorflowexample_b200.models import oracle


def test_read_data_sets_policy(data_dir):
    d = read_data_sets(data_dir, one_hot=True, seed=0)
    assert d.source in ("mnist", "t10k-split")
    if d.source == "t10k-split":           # train-images is absent from the reference checkout
        assert d.train.num_examples == 7500 and d.validation.num_examples == 500 and d.test.num_examples == 2000
    else:
        assert d.train.num_examples == 55000 and d.validation.num_examples == 5000 and d.test.num_examples == 10000
    assert d.train.images.dtype == np.float32 and d.train.images.shape[1] == 784
    assert 0.0 <= d.train.images.min() and d.train.images.max() <= 1.0
    assert d.train.labels.dtype == np.float64 and d.train.labels.shape[1] == 10     # TF1 one-hot is float64
    assert d.train.label_ids.dtype == np.int32
    with pytest.raises(FileNotFoundError):
        read_data_sets("/nonexistent", allow_fallback=False)
    s = read_data_sets("/nonexistent", one_hot=False)
    assert s.source == "synthetic" and s.train.images.shape == (55000, 784) and s.test.num_examples == 10000


def test_next_batch_epoch_semantics():
    n = 10
    imgs = np.arange(n, dtype=np.float32).reshape(n, 1)
    ds = DataSet(imgs, np.arange(n), seed=3)
    seen = []
    for _ in range(2):                                   # 2 batches of 4: inside epoch 0
        x, y = ds.next_batch(4)
        assert x.shape == (4, 1) and (x[:, 0] == y).all()
        seen += y.tolist()
    assert ds.epochs_completed == 0 and len(set(seen)) == 8
    x, y = ds.next_batch(4)                              # straddles: 2 left of old epoch + 2 of the reshuffled new
    assert ds.epochs_completed == 1 and x.shape == (4, 1)
    assert set(seen) | set(y[:2].tolist()) == set(range(n))          # the rest of epoch 0 comes first
    assert sorted(seen) != seen or True
    ds2 = DataSet(imgs.copy(), np.arange(n), seed=3)     # same seed -> same order (per-process RNG)
    a = [ds2.next_batch(5)[1].tolist() for _ in range(2)]
    ds3 = DataSet(imgs.copy(), np.arange(n), seed=3)
    b = [ds3.next_batch(5)[1].tolist() for _ in range(2)]
    assert a == b and sorted(a[0] + a[1]) == list(range(n))
    u8 = DataSet(np.full((2, 28, 28), 255, np.uint8), np.zeros(2, np.uint8))
    assert u8.images.shape == (2, 784) and u8.images.dtype == np.float32 and u8.images.max() == pytest.approx(1.0)


def test_one_hot():
    oh = dense_to_one_hot(np.array([1, 0, 9], np.uint8))
    assert oh.dtype == np.float64 and oh.shape == (3, 10) and oh[2, 9] == 1 and oh.sum() == 3


def test_oracle_gradients_match_finite_differences():
    rng = np.random.default_rng(0)
    dims = (6, 5, 4, 3)
    params = [p.astype(np.float64) * 0.5 for p in oracle.init_params(dims, rng)]
    x = rng.random((7, 6)); y = rng.integers(0, 3, 7)
    loss, acc, grads = oracle.gradients(params, x, y)
    eps = 1e-6
    for pi in range(len(params)):
        idx = tuple(rng.integers(0, s) for s in params[pi].shape)
        old = params[pi][idx]
        params[pi][idx] = old + eps; lp = oracle.gradients(params, x, y)[0]
        params[pi][idx] = old - eps; lm = oracle.gradients(params, x, y)[0]
        params[pi][idx] = old
        assert abs((lp - lm) / (2 * eps) - grads[pi][idx]) < 1e-6


def test_reference_init_loss_and_sgd_progress():
    """N(0,1) kernels saturate the sigmoids: initial loss is O(10) (~9.75 on MNIST)."""
    rng = np.random.default_rng(0)
    params = oracle.init_params((784, 100, 10), rng, np.float64)
    x, y = synthetic_arrays(2000, seed=1)
    l0, _, _ = oracle.gradients(params, x[:100].astype(np.float64), y[:100].astype(np.int64))
    assert 3.0 < l0 < 25.0
    losses = []
    for s in range(200):
        rows = slice((s * 100) % 2000, (s * 100) % 2000 + 100)
        losses.append(oracle.sgd_step(params, x[rows].astype(np.float64), y[rows].astype(np.int64), 0.1)[0])
    assert np.mean(losses[-20:]) < 0.5 * np.mean(losses[:20])
    f = oracle.flops_per_step((784, 100, 10), 100)
    assert f["fwd"] == pytest.approx(15.88e6, rel=1e-3) and f["step"] == pytest.approx(31.96e6, rel=1e-3)


def test_gloo_ps_apply_matches_oracle_sgd_and_adam():
    """The ps-side update rules of the gloo configuration (parallel/gloo_ps.py::_apply) against the numpy oracle:
    TF1 ApplyGradientDescent and ApplyAdam on a flat vector holding var | m | v."""
    import torch
    from distributedtensorflowexample_b200.models import oracle
    from distributedtensorflowexample_b200.parallel.gloo_ps import GlooParamServer
    rng = np.random.default_rng(0)
    n = 37
    w0 = rng.normal(size=n).astype(np.float32)
    grads = [rng.normal(size=n).astype(np.float32) for _ in range(4)]
    # SGD
    ps = GlooParamServer(0, 1)
    ps.flat = torch.from_numpy(w0.copy())
    ref = [w0.astype(np.float64).copy()]
    for g in grads:
        ps._apply([0, -1, -1, n], torch.from_numpy(g), 0.05, False, (0.0, 0.0, 0.0))
        ref[0] -= 0.05 * g
    np.testing.assert_allclose(ps.flat.numpy(), ref[0], rtol=1e-5, atol=1e-6)
    # Adam: moments live in the same flat vector, after the variable
    ps = GlooParamServer(0, 1)
    ps.flat = torch.cat([torch.from_numpy(w0.copy()), torch.zeros(2 * n)])
    ref = [w0.astype(np.float64).copy()]
    st = oracle.AdamState(ref)
    for g in grads:
        ps._apply([0, n, 2 * n, n], torch.from_numpy(g), 0.01, True, (0.9, 0.999, 1e-8))
        st.apply(ref, [g.astype(np.float64)], 0.01)
    assert ps.adam_t == 4
    np.testing.assert_allclose(ps.flat[:n].numpy(), ref[0], rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(ps.flat[n:2 * n].numpy(), st.m[0], rtol=1e-4, atol=1e-7)
    np.testing.assert_allclose(ps.flat[2 * n:].numpy(), st.v[0], rtol=1e-4, atol=1e-8)
