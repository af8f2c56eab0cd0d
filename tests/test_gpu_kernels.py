"""GPU numerics: every fp32 CUDA kernel against a plain PyTorch/numpy fp32 restatement."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from distributedtensorflowexample_b200.models import oracle


def _ops():
    from distributedtensorflowexample_b200.ops import cuda_ops
    return cuda_ops


@pytest.mark.parametrize("M,K,N,act", [(100, 784, 100, 1), (100, 100, 10, 0), (257, 300, 130, 1), (10000, 784, 100, 1)])
def test_dense_fwd(M, K, N, act):
    ops = _ops()
    g = torch.Generator(device="cuda").manual_seed(1)
    x = torch.rand(M, K, device="cuda", generator=g)
    W = torch.randn(K, N, device="cuda", generator=g) * 0.1
    b = torch.randn(N, device="cuda", generator=g)
    out = ops.dense_fwd(x, W, b, act)
    ref = x.double() @ W.double() + b.double()
    if act:
        ref = torch.sigmoid(ref)
    torch.testing.assert_close(out.double(), ref, rtol=1e-4, atol=1e-4)


@pytest.mark.parametrize("M,K,N", [(100, 100, 10), (100, 784, 100), (300, 257, 65)])
def test_dense_bwd(M, K, N):
    ops = _ops()
    g = torch.Generator(device="cuda").manual_seed(2)
    X = torch.rand(M, K, device="cuda", generator=g)           # layer input (sigmoid output of prev layer)
    dY = torch.randn(M, N, device="cuda", generator=g) * 0.1
    W = torch.randn(K, N, device="cuda", generator=g)
    dX = ops.dense_bwd_dx(dY, W, X, ops.ACT_SIGMOID)
    ref_dx = (dY.double() @ W.double().T) * X.double() * (1 - X.double())
    torch.testing.assert_close(dX.double(), ref_dx, rtol=1e-4, atol=1e-5)
    W0 = W.clone()
    G = torch.empty_like(W)
    ops.dense_bwd_dw_push(X, dY, W, -0.5, atomic=True, grad_out=G)
    ref_g = X.double().T @ dY.double()
    torch.testing.assert_close(G.double(), ref_g, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(W.double(), W0.double() - 0.5 * ref_g, rtol=1e-4, atol=1e-4)
    b = torch.zeros(N, device="cuda")
    ops.bias_grad_push(dY, b, -1.0, atomic=False)
    torch.testing.assert_close(b.double(), -dY.double().sum(0), rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("M,C", [(100, 10), (1000, 10), (64, 40)])
def test_softmax_xent(M, C):
    ops = _ops()
    rng = np.random.default_rng(3)
    z = rng.normal(0, 3, size=(M, C)).astype(np.float32)
    y = rng.integers(0, C, size=M).astype(np.int32)
    loss, dz, rowloss = oracle.softmax_xent(z.astype(np.float64), y)
    zt = torch.tensor(z, device="cuda"); yt = torch.tensor(y, device="cuda")
    dZ = torch.empty(M, C, device="cuda"); rl = torch.empty(M, device="cuda")
    rc = torch.empty(M, device="cuda", dtype=torch.int32)
    ops.softmax_xent(zt, yt, 1.0 / M, dZ, rl, rc)
    out_l = torch.zeros(1, device="cuda"); out_a = torch.zeros(1, device="cuda")
    ops.mean_loss_acc(rl, rc, M, out_l, out_a)
    np.testing.assert_allclose(dZ.cpu().numpy(), dz, rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(rl.cpu().numpy(), rowloss, rtol=1e-4, atol=1e-5)
    assert abs(out_l.item() - loss) < 1e-4 * max(1, abs(loss))
    assert abs(out_a.item() - oracle.accuracy(z, y)) < 1e-6
    # dense one-hot labels (the reference's feed) give the same answer
    onehot = torch.zeros(M, C, device="cuda"); onehot[torch.arange(M), yt.long()] = 1
    dZ2 = torch.empty_like(dZ); rl2 = torch.empty_like(rl); rc2 = torch.empty_like(rc)
    ops.softmax_xent(zt, onehot, 1.0 / M, dZ2, rl2, rc2)
    torch.testing.assert_close(dZ2, dZ, rtol=1e-5, atol=1e-7)
    torch.testing.assert_close(rl2, rl, rtol=1e-5, atol=1e-6)
    assert torch.equal(rc2, rc)


@pytest.mark.parametrize("n", [10, 4096, 79510, 1 << 20])
def test_push_pull_flat(n):
    ops = _ops()
    shard = torch.randn(n + 8, device="cuda")[:n] if False else torch.randn(n, device="cuda")
    grad = torch.randn(n, device="cuda")
    ref = shard.double() - 0.25 * grad.double()
    ops.push_sgd(shard, grad, -0.25, atomic=True)
    torch.testing.assert_close(shard.double(), ref, rtol=1e-6, atol=1e-6)
    dst = torch.empty(n, device="cuda")
    ops.pull(dst, shard)
    assert torch.equal(dst, shard)
    ops.push_sgd(shard, grad, 0.25, atomic=False)
    torch.testing.assert_close(shard.double(), ref + 0.25 * grad.double(), rtol=1e-6, atol=1e-6)


def test_push_bf16():
    ops = _ops()
    n = 4096 + 6
    shard = torch.randn(n, device="cuda").bfloat16()
    grad = torch.randn(n, device="cuda")
    ref = shard.float() + 0.5 * grad
    ops.push_sgd(shard, grad, 0.5)
    torch.testing.assert_close(shard.float(), ref, rtol=2e-2, atol=2e-2)


def test_adam_apply():
    ops = _ops()
    n = 5000
    rng = np.random.default_rng(0)
    p = rng.normal(size=n).astype(np.float32)
    params = [p.copy()]
    st = oracle.AdamState(params)
    P = torch.tensor(p, device="cuda"); M = torch.zeros(n, device="cuda"); V = torch.zeros(n, device="cuda")
    slot = torch.zeros(n, device="cuda")
    for t in range(1, 4):
        g = rng.normal(size=n).astype(np.float32)
        st.apply(params, [g], 1e-2)
        ops.push_sgd(slot, torch.tensor(g, device="cuda"), 1.0, atomic=True)   # workers accumulate into the slot
        lr_t = 1e-2 * np.sqrt(1 - 0.999 ** t) / (1 - 0.9 ** t)
        ops.adam_apply(P, M, V, slot, lr_t, 0.9, 0.999, 1e-8)
        assert float(slot.abs().max()) == 0.0
    np.testing.assert_allclose(P.cpu().numpy(), params[0], rtol=1e-4, atol=1e-5)


def test_global_step_add():
    ops = _ops()
    gs = torch.zeros(1, device="cuda", dtype=torch.int32)
    obs = torch.full((1,), -1, device="cuda", dtype=torch.int32)
    for i in range(5):
        ops.global_step_add(gs, 1, obs)
        assert obs.item() == i
    assert gs.item() == 5


def test_sync_rs_apply_single_rank():
    ops = _ops()
    n = 79510
    grads = torch.randn(n, device="cuda")
    param = torch.randn(n, device="cuda")
    ref = param.double() - 0.1 * grads.double()
    flags = torch.zeros(2 * ops.lib().dtfe_sync_max_ranks(), device="cuda", dtype=torch.int32)
    local_sync = torch.zeros(4, device="cuda", dtype=torch.int32)
    gs = torch.zeros(1, device="cuda", dtype=torch.int32)
    ops.sync_rs_apply([grads.data_ptr()], [flags.data_ptr()], 0, 1, [(param[:1000], 0, 1000), (param[1000:], 1000, n)], -0.1, 1, local_sync, gs)
    torch.testing.assert_close(param.double(), ref, rtol=1e-6, atol=1e-6)
    assert gs.item() == 1
    ops.sync_rs_apply([grads.data_ptr()], [flags.data_ptr()], 0, 1, [(param, 0, n)], -0.1, 2, local_sync, gs)
    torch.testing.assert_close(param.double(), ref - 0.1 * grads.double(), rtol=1e-6, atol=1e-5)
    assert gs.item() == 2 and local_sync[2].item() == 0
    # a peer that never arrives: the wait is abandoned after the timeout instead of hanging the GPU
    flags2 = torch.zeros_like(flags)
    before = param.clone()
    ops.sync_rs_apply([grads.data_ptr(), grads.data_ptr()], [flags2.data_ptr(), flags2.data_ptr()], 0, 2,
                      [(param, 0, n)], -0.1, 1, torch.zeros(4, device="cuda", dtype=torch.int32), None, timeout_s=0.05)
    torch.cuda.synchronize()
    assert torch.equal(param, before)


def test_contended_push_lossless_vs_unlocked():
    """Two streams push into the same shard at the same time (what two workers sharing a ps do).  The default
    push (red.add, performed by the L2 that owns the address) must conserve every update exactly - the values are
    powers of two, so fp32 addition is exact in any order.  --unlocked_apply (TF1 use_locking=False: load, add,
    store) may lose updates under contention and must never invent any; the fraction lost is printed."""
    from distributedtensorflowexample_b200.ops import cuda_ops as ops
    n, reps = 1 << 20, 200
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    g1 = torch.full((n,), 0.25, device="cuda"); g2 = torch.full((n,), 0.5, device="cuda")
    expect = -reps * (0.25 + 0.5)
    lost = {}
    for atomic in (True, False):
        shard = torch.zeros(n, device="cuda")
        torch.cuda.synchronize()
        for s, g in ((s1, g1), (s2, g2)):
            with torch.cuda.stream(s):
                for _ in range(reps):
                    ops.push_sgd(shard, g, -1.0, atomic=atomic)
        torch.cuda.synchronize()
        got = shard.cpu().numpy()
        if atomic:
            assert np.all(got == expect), (got.min(), got.max(), expect)
        else:
            assert np.all(got <= 0) and np.all(got >= expect)           # nothing invented ...
            assert np.all(np.abs(np.round(got * 4) - got * 4) < 1e-6)   # ... every value is a sum of whole updates
            lost[atomic] = float(1.0 - got.sum() / (expect * n))
    print(f"unlocked apply under contention lost {100 * lost[False]:.2f} % of the updates; red.add lost none")
