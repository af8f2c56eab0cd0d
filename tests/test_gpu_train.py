"""GPU end-to-end parity: fused and layer-wise step engines vs the numpy oracle (same init, same batches)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from distributedtensorflowexample_b200.models import oracle
from distributedtensorflowexample_b200.models.mlp import MLPSpec, reference_mlp_spec
from distributedtensorflowexample_b200.train.replica import build_replicated_model


def _store(spec, num_ps=1, seed=0):
    from distributedtensorflowexample_b200.parallel.ps import PeerParamStore
    model = build_replicated_model(spec, num_ps=num_ps)
    store = PeerParamStore(model.global_all, num_ps, torch.device("cuda:0"))
    store.allocate()
    store.initialize(np.random.default_rng(seed))
    names = [v.name for v in model.global_trainable]
    return model, store, names


def _data(n=1000, seed=5):
    from distributedtensorflowexample_b200.data import synthetic_arrays
    x, y = synthetic_arrays(n, seed=seed)
    return x, y.astype(np.int32)


@pytest.mark.parametrize("num_ps", [1, 3])
def test_store_layout_and_init(num_ps):
    spec = reference_mlp_spec()
    model, store, names = _store(spec, num_ps)
    assert names == ["global/dense/kernel", "global/dense/bias", "global/dense_1/kernel", "global/dense_1/bias"]
    assert store.tensor("global/dense/kernel").shape == (784, 100)
    assert store.read("global/global_step").item() == 0
    assert float(np.abs(store.read("global/dense/bias")).max()) == 0.0
    w = store.read("global/dense/kernel")
    assert 0.9 < w.std() < 1.1
    assert not store.is_ready()
    store.mark_ready()
    assert store.is_ready()
    # round-robin placement over ps tasks in creation order
    assert [store.shard_of(n) for n in names] == [i % num_ps for i in range(4)]
    store.close()


@pytest.mark.parametrize("ks", [8, 16])
@pytest.mark.parametrize("atomic", [True, False])
def test_fused_step_matches_oracle(ks, atomic):
    from distributedtensorflowexample_b200.train.gpu_step import FusedMlp2Stepper
    spec = reference_mlp_spec()
    model, store, names = _store(spec)
    params_np = [store.read(n).astype(np.float64) for n in names]
    x, y = _data()
    X = torch.tensor(x, device="cuda"); Y = torch.tensor(y, device="cuda")
    B, lr, nsteps = 100, 0.05, 6
    st = FusedMlp2Stepper(spec, B, [store.tensor(n) for n in names], store.tensor("global/global_step"), lr,
                          "cuda:0", max_steps_per_launch=nsteps, atomic=atomic, ks=ks)
    rng = np.random.default_rng(1)
    idx = rng.integers(0, x.shape[0], size=nsteps * B).astype(np.int32)
    st.run(X, Y, nsteps=nsteps, batch_idx=torch.tensor(idx, device="cuda"))
    torch.cuda.synchronize()
    losses, accs = [], []
    for s in range(nsteps):
        rows = idx[s * B:(s + 1) * B]
        l, a = oracle.sgd_step(params_np, x[rows].astype(np.float64), y[rows], lr)
        losses.append(l); accs.append(a)
    np.testing.assert_allclose(st.loss[:nsteps].cpu().numpy(), losses, rtol=2e-4, atol=2e-4)
    np.testing.assert_allclose(st.acc[:nsteps].cpu().numpy(), accs, atol=1e-6)
    assert st.step[:nsteps].cpu().tolist() == list(range(nsteps))
    assert store.read("global/global_step").item() == nsteps
    for n, p in zip(names, params_np):
        np.testing.assert_allclose(store.read(n), p, rtol=1e-3, atol=2e-4)
    # a second launch must start from clean sync words
    st.run(X, Y, nsteps=2, batch_idx=torch.tensor(idx[:2 * B], device="cuda"))
    torch.cuda.synchronize()
    assert store.read("global/global_step").item() == nsteps + 2
    store.close()


@pytest.mark.parametrize("dims,B,ks", [((50, 30, 7), 33, 8), ((100, 128, 16), 128, 8), ((130, 20, 3), 5, 16),
                                       ((64, 30, 7), 33, 8), ((96, 64, 10), 50, 8),
                                       # 138 W1 slices leave 10 helper CTAs: too few for H=64, so every CTA takes small
                                       # gradients after its dW1, and slice CTAs double as fetchers of the small parameters
                                       ((1100, 64, 10), 32, 8)])
def test_fused_step_odd_shapes_match_oracle(dims, B, ks):
    """Shapes that are not whole float4s / whole K slices: partial cp.async quads, the scalar split-K
    reduction, a short last W1 slice, grids with and without helper CTAs."""
    from distributedtensorflowexample_b200.data import synthetic_arrays
    from distributedtensorflowexample_b200.train.gpu_step import FusedMlp2Stepper
    spec = MLPSpec(dims)
    model, store, names = _store(spec, seed=2)
    params_np = [store.read(n).astype(np.float64) for n in names]
    x, y = synthetic_arrays(400, seed=7, num_classes=dims[-1], dim=dims[0])
    y = y.astype(np.int32)
    X = torch.tensor(x, device="cuda"); Y = torch.tensor(y, device="cuda")
    lr, nsteps = 0.05, 5
    st = FusedMlp2Stepper(spec, B, [store.tensor(n) for n in names], store.tensor("global/global_step"), lr,
                          "cuda:0", max_steps_per_launch=nsteps, ks=ks)
    idx = np.random.default_rng(4).integers(0, x.shape[0], size=nsteps * B).astype(np.int32)
    st.run(X, Y, nsteps=nsteps, batch_idx=torch.tensor(idx, device="cuda"))
    torch.cuda.synchronize()
    losses = []
    for s in range(nsteps):
        rows = idx[s * B:(s + 1) * B]
        losses.append(oracle.sgd_step(params_np, x[rows].astype(np.float64), y[rows], lr)[0])
    np.testing.assert_allclose(st.loss[:nsteps].cpu().numpy(), losses, rtol=2e-4, atol=2e-4)
    assert st.step[:nsteps].cpu().tolist() == list(range(nsteps))
    for n, p in zip(names, params_np):
        np.testing.assert_allclose(store.read(n), p, rtol=1e-3, atol=2e-4)
    store.close()


_PULL_MODE_SCRIPT = r"""
import sys, numpy as np, torch
sys.path.insert(0, {root!r})
from distributedtensorflowexample_b200.models.mlp import reference_mlp_spec
from distributedtensorflowexample_b200.train.replica import build_replicated_model
from distributedtensorflowexample_b200.parallel.ps import PeerParamStore
from distributedtensorflowexample_b200.train.gpu_step import FusedMlp2Stepper
from distributedtensorflowexample_b200.data import synthetic_arrays
spec = reference_mlp_spec()
model = build_replicated_model(spec, num_ps=1)
store = PeerParamStore(model.global_all, 1, torch.device("cuda:0")); store.allocate()
store.initialize(np.random.default_rng(0))
names = [v.name for v in model.global_trainable]
x, y = synthetic_arrays(1000, seed=5)
X = torch.tensor(x, device="cuda"); Y = torch.tensor(y.astype(np.int32), device="cuda")
st = FusedMlp2Stepper(spec, 100, [store.tensor(n) for n in names], store.tensor("global/global_step"), 0.05,
                      "cuda:0", max_steps_per_launch=12)
idx = np.random.default_rng(1).integers(0, 1000, size=1200).astype(np.int32)
st.run(X, Y, nsteps=12, batch_idx=torch.tensor(idx, device="cuda"))
st.run(X, Y, nsteps=3, batch_idx=torch.tensor(idx[:300], device="cuda"))
torch.cuda.synchronize()
np.savez({out!r}, loss=st.loss[:3].cpu().numpy(), **{{n.replace("/", "."): store.read(n) for n in names}})
"""


def test_fused_pull_ahead_is_bitwise_the_direct_pull(tmp_path):
    """With one worker the pull that runs a step ahead (fetch before the push + own update added locally)
    must give exactly the bits of the pull that re-reads the shard after the push."""
    import os, subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    outs = []
    for mode in ("ahead", "direct"):
        out = str(tmp_path / f"{mode}.npz")
        env = dict(os.environ)
        env.pop("DTFE_MLP2_NO_AHEAD", None)
        if mode == "direct":
            env["DTFE_MLP2_NO_AHEAD"] = "1"
        subprocess.run([sys.executable, "-c", _PULL_MODE_SCRIPT.format(root=root, out=out)], check=True, env=env,
                       timeout=300)
        outs.append(np.load(out))
    assert sorted(outs[0].files) == sorted(outs[1].files) and len(outs[0].files) == 5
    for k in outs[0].files:
        np.testing.assert_array_equal(outs[0][k], outs[1][k], err_msg=k)


def test_fused_single_step_contiguous_batch_and_eval_mode():
    from distributedtensorflowexample_b200.train.gpu_step import FusedMlp2Stepper
    spec = reference_mlp_spec()
    model, store, names = _store(spec, seed=3)
    params_np = [store.read(n).astype(np.float64) for n in names]
    x, y = _data(100)
    X = torch.tensor(x, device="cuda"); Y = torch.tensor(y, device="cuda")
    st = FusedMlp2Stepper(spec, 100, [store.tensor(n) for n in names], store.tensor("global/global_step"), 0.001, "cuda:0")
    st.run(X, Y, nsteps=1, train=False)
    torch.cuda.synchronize()
    loss, acc, _ = oracle.gradients(params_np, x.astype(np.float64), y)
    assert abs(st.loss[0].item() - loss) < 2e-4 * max(1.0, loss)
    assert abs(st.acc[0].item() - acc) < 1e-6
    assert store.read("global/global_step").item() == 0
    for n, p in zip(names, params_np):        # eval mode must not touch the shard
        np.testing.assert_array_equal(store.read(n).astype(np.float64), p)
    # initial loss with the reference init is ~9.75 on MNIST-like data: sanity band
    assert 2.0 < loss < 30.0
    store.close()


@pytest.mark.parametrize("dims,B", [((784, 100, 10), 100), ((784, 64, 48, 10), 37)])
def test_layerwise_step_matches_oracle(dims, B):
    from distributedtensorflowexample_b200.train.gpu_step import LayerwiseStepper
    spec = MLPSpec(dims)
    model, store, names = _store(spec, num_ps=2)
    params_np = [store.read(n).astype(np.float64) for n in names]
    x, y = _data(400)
    X = torch.tensor(x, device="cuda"); Y = torch.tensor(y, device="cuda")
    lr = 0.05
    st = LayerwiseStepper(spec, B, [store.tensor(n) for n in names], store.tensor("global/global_step"), lr, "cuda:0")
    for s in range(4):
        rows = slice(s * B, (s + 1) * B)
        st.run(X[rows], Y[rows])
        l, a = oracle.sgd_step(params_np, x[rows].astype(np.float64), y[rows], lr)
        assert abs(st.loss.item() - l) < 2e-4 * max(1.0, l)
        assert abs(st.acc.item() - a) < 1e-6
        assert st.step.item() == s
    for n, p in zip(names, params_np):
        np.testing.assert_allclose(store.read(n), p, rtol=1e-3, atol=2e-4)
    correct, total = st.eval_counts(X, Y)
    logits, _ = oracle.forward(params_np, x.astype(np.float64))
    assert total == 400
    assert abs(correct.item() - int((logits.argmax(1) == y).sum())) <= 1
    store.close()


def test_trainer_device_and_host_paths_agree():
    from distributedtensorflowexample_b200.train.gpu_step import DeviceDataset
    from distributedtensorflowexample_b200.train.trainer import GpuTrainer
    spec = reference_mlp_spec()
    x, y = _data(2000)
    outs = []
    for mode in ("device", "host", "host_native"):
        model, store, names = _store(spec, seed=11)
        tr = GpuTrainer(spec, 100, 0.01, [store.tensor(n) for n in names], store.tensor("global/global_step"),
                        "cuda:0", steps_per_launch=5)
        ds = DeviceDataset(x, y, "cuda:0", seed=7)
        if mode == "device":
            tr.train_device_steps(ds, 10)
        elif mode == "host_native":
            xs, ys = [], []
            for s in range(10):
                rows = ds.next_indices(100, 1).cpu().long()
                xs.append(torch.tensor(x)[rows].contiguous().pin_memory()); ys.append(torch.tensor(y)[rows].contiguous().pin_memory())
            losses = torch.zeros(10).pin_memory()
            tr.train_host_batches(xs, ys, losses)
            torch.cuda.synchronize()
            assert torch.isfinite(losses).all() and losses[0] > 0
        else:
            xp = torch.tensor(x).pin_memory(); yp = torch.tensor(y).pin_memory()
            for s in range(10):
                rows = ds.next_indices(100, 1).cpu().long()
                xb = xp.index_select(0, rows).pin_memory(); yb = yp.index_select(0, rows).pin_memory()
                tr.train_host_batch(xb, yb, out_index=s)
                torch.cuda.synchronize()
        torch.cuda.synchronize()
        outs.append([store.read(n) for n in names] + [store.read("global/global_step")])
        store.close()
    for other in outs[1:]:
        for a, b in zip(outs[0], other):
            np.testing.assert_allclose(a, b, rtol=1e-5, atol=1e-6)
    assert outs[0][-1].item() == 10


@pytest.mark.parametrize("engine", ["fused", "layerwise"])
def test_pipelined_launches_equal_the_blocking_path(engine):
    """launch_async/collect (what Worker.learn drives): two launch groups in flight give the same parameters and the
    same per-step metrics as the blocking calls, for every step of every group."""
    from distributedtensorflowexample_b200.train.gpu_step import DeviceDataset
    from distributedtensorflowexample_b200.train.trainer import GpuTrainer
    spec = reference_mlp_spec()
    x, y = _data(3000)
    res = []
    for mode in ("blocking", "pipelined"):
        model, store, names = _store(spec, seed=3)
        tr = GpuTrainer(spec, 100, 0.02, [store.tensor(n) for n in names], store.tensor("global/global_step"),
                        "cuda:0", steps_per_launch=1 if mode == "blocking" else 8, engine=engine)
        ds = DeviceDataset(x, y, "cuda:0", seed=9)
        losses, steps = [], []
        if mode == "blocking":
            for _ in range(19):
                l, a, s = tr.train_device_steps(ds, 1)
                losses.append(float(l[0].item())); steps.append(int(s[0].item()))
        else:
            sizes, tickets = [8, 8, 3], []
            for n in sizes:
                tickets.append(tr.launch_async(ds, n, ring_depth=3, ring_cap=8))
            for tk in tickets:
                l, a, s = tr.collect(tk)
                assert len(l) == tk.nsteps
                losses += l.tolist(); steps += s.tolist()
        torch.cuda.synchronize()
        res.append((losses, steps, [store.read(n) for n in names]))
        store.close()
    assert res[0][1] == res[1][1] == list(range(19))
    np.testing.assert_allclose(res[0][0], res[1][0], rtol=1e-5, atol=1e-6)
    for a, b in zip(res[0][2], res[1][2]):
        np.testing.assert_allclose(a, b, rtol=1e-5, atol=1e-6)


def test_layerwise_adam_matches_oracle():
    """Optional Adam (north star): workers push raw gradients into owner-side slots, the owner applies."""
    from distributedtensorflowexample_b200.parallel.ps import PeerParamStore
    from distributedtensorflowexample_b200.train.gpu_step import LayerwiseStepper
    spec = reference_mlp_spec()
    model = build_replicated_model(spec, num_ps=2, optimizer="adam")
    store = PeerParamStore(model.global_all, 2, torch.device("cuda:0"))
    store.allocate()
    store.initialize(np.random.default_rng(5))
    names = [v.name for v in model.global_trainable]
    assert "global/dense/kernel/Adam" in store.layout.entries and "global/dense_1/bias/Adam_1" in store.layout.entries
    assert store.shard_of("global/dense/kernel/Adam") == store.shard_of("global/dense/kernel")
    params_np = [store.read(n).astype(np.float64) for n in names]
    st_ref = oracle.AdamState(params_np)
    x, y = _data(500)
    X = torch.tensor(x, device="cuda"); Y = torch.tensor(y, device="cuda")
    adam = dict(slots=[store.tensor(n + "/grad_slot") for n in names], m=[store.tensor(n + "/Adam") for n in names],
                v=[store.tensor(n + "/Adam_1") for n in names], apply=True)
    B, lr = 100, 0.01
    st = LayerwiseStepper(spec, B, [store.tensor(n) for n in names], store.tensor("global/global_step"), lr, "cuda:0",
                          adam=adam)
    for s in range(4):
        rows = slice(s * B, (s + 1) * B)
        st.run(X[rows], Y[rows])
        _, _, grads = oracle.gradients(params_np, x[rows].astype(np.float64), y[rows])
        st_ref.apply(params_np, grads, lr)
    for n, p in zip(names, params_np):
        np.testing.assert_allclose(store.read(n), p, rtol=2e-3, atol=5e-4)
    assert store.read("global/global_step").item() == 4
    assert float(np.abs(store.read(names[0] + "/grad_slot")).max()) == 0.0
    store.close()


def test_layerwise_tf32_tensor_core_path():
    """--tf32: layer 1 of the reference model (fp32 [784,100] kernel in the shard) runs on tcgen05 kind::tf32 with
    the weights pulled by TMA; the 10-wide output layer stays on the fp32 kernels."""
    from distributedtensorflowexample_b200.train.gpu_step import LayerwiseStepper
    spec = reference_mlp_spec()
    model, store, names = _store(spec, num_ps=1, seed=21)
    for n in names:                                   # unsaturated sigmoids so that tf32 rounding is visible
        if store.tensor(n).dim() == 2:
            store.assign(n, (np.random.default_rng(1).normal(size=store.tensor(n).shape) * 0.05).astype(np.float32))
    params_np = [store.read(n).astype(np.float64) for n in names]
    x, y = _data(400)
    X = torch.tensor(x, device="cuda"); Y = torch.tensor(y, device="cuda")
    B, lr = 100, 0.5
    st = LayerwiseStepper(spec, B, [store.tensor(n) for n in names], store.tensor("global/global_step"), lr, "cuda:0",
                          tf32=True)
    for s in range(3):
        rows = slice(s * B, (s + 1) * B)
        st.run(X[rows], Y[rows])
        l, a = oracle.sgd_step(params_np, x[rows].astype(np.float64), y[rows], lr)
        assert abs(st.loss.item() - l) < 5e-3 * max(1.0, l)
    for n, p in zip(names, params_np):
        np.testing.assert_allclose(store.read(n), p, rtol=0, atol=3e-3 * (np.abs(p).max() + 1))
    assert store.read("global/global_step").item() == 3
    store.close()


def test_fused_step_with_in_kernel_sync_exchange_single_rank():
    """Synchronous mode with the exchange as phase D of the step kernel (world = 1: the protocol still runs -
    flags, epochs, reduce of the single buffer, apply): must equal plain SGD, several steps per launch."""
    from distributedtensorflowexample_b200.parallel.sync import SyncGroup
    from distributedtensorflowexample_b200.train.gpu_step import FusedMlp2Stepper
    spec = reference_mlp_spec()
    model, store, names = _store(spec, num_ps=2, seed=13)
    params_np = [store.read(n).astype(np.float64) for n in names]
    x, y = _data(1000)
    X = torch.tensor(x, device="cuda"); Y = torch.tensor(y, device="cuda")
    B, lr, nsteps = 100, 0.05, 5
    sg = SyncGroup(store, names, 0, 1, timeout_s=5.0)
    shapes = {n: tuple(store.tensor(n).shape) for n in names}
    st = FusedMlp2Stepper(spec, B, [store.tensor(n) for n in names], store.tensor("global/global_step"), lr, "cuda:0",
                          max_steps_per_launch=nsteps, grad_buffers=sg.grad_views(shapes))
    st.sync_exchange = sg.exchange_struct(lr)
    idx = torch.arange(nsteps * B, device="cuda", dtype=torch.int32)
    st.run(X, Y, nsteps=nsteps, batch_idx=idx)
    torch.cuda.synchronize()
    assert sg.aborted() == 0
    for s in range(nsteps):
        rows = slice(s * B, (s + 1) * B)
        oracle.sgd_step(params_np, x[rows].astype(np.float64), y[rows], lr)
    for n, p in zip(names, params_np):
        np.testing.assert_allclose(store.read(n), p, rtol=1e-3, atol=2e-4)
    assert store.read("global/global_step").item() == nsteps
    assert st.step[:nsteps].cpu().tolist() == list(range(nsteps))
    # the separate exchange kernel continues the same epoch sequence
    st2 = FusedMlp2Stepper(spec, B, [store.tensor(n) for n in names], store.tensor("global/global_step"), lr, "cuda:0",
                           grad_buffers=sg.grad_views(shapes))
    st2.run(X, Y, nsteps=1, batch_idx=idx[:B])
    sg.reduce_scatter_apply(lr)
    torch.cuda.synchronize()
    assert store.read("global/global_step").item() == nsteps + 1 and sg.aborted() == 0
    sg.close(); store.close()


@pytest.mark.parametrize("mode", ["allreduce", "allreduce_ll"])
@pytest.mark.parametrize("dims,B", [((784, 100, 10), 100), ((50, 30, 7), 37)])
def test_fused_step_allreduce_exchange_single_rank(dims, B, mode):
    """The all-reduce form of the in-kernel exchange with world = 1: per-CTA flags, alternating gradient copies,
    W1 slice resident in shared memory across steps and launches, apply to the home copy, local step counter.
    Must equal plain SGD over two launches (the second starts from what the first wrote back)."""
    from distributedtensorflowexample_b200.parallel.sync import SyncGroup
    from distributedtensorflowexample_b200.train.gpu_step import FusedMlp2Stepper
    spec = MLPSpec(dims)
    model, store, names = _store(spec, num_ps=2, seed=21)
    params_np = [store.read(n).astype(np.float64) for n in names]
    from distributedtensorflowexample_b200.data import synthetic_arrays
    x, y = synthetic_arrays(1000, seed=5, num_classes=dims[2], dim=dims[0])
    y = y.astype(np.int32)
    X = torch.tensor(x, device="cuda"); Y = torch.tensor(y, device="cuda")
    lr, nsteps = 0.05, 4
    sg = SyncGroup(store, names, 0, 1, timeout_s=5.0, mode=mode)
    sg.refresh_replicas()
    shapes = {n: tuple(store.tensor(n).shape) for n in names}
    st = FusedMlp2Stepper(spec, B, sg.home_params(), sg.home("global/global_step"), lr, "cuda:0",
                          max_steps_per_launch=nsteps, remote=False, grad_buffers=sg.grad_views(shapes))
    st.sync_exchange = sg.exchange_struct(lr)
    idx = torch.arange(2 * nsteps * B, device="cuda", dtype=torch.int32)
    st.run(X, Y, nsteps=nsteps, batch_idx=idx[:nsteps * B])
    st.run(X, Y, nsteps=nsteps, batch_idx=idx[nsteps * B:], out_offset=0)
    torch.cuda.synchronize()
    assert sg.aborted() == 0
    for s in range(2 * nsteps):
        rows = slice(s * B, (s + 1) * B)
        oracle.sgd_step(params_np, x[rows].astype(np.float64), y[rows], lr)
    for n, p in zip(names, params_np):
        np.testing.assert_allclose(store.read(n), p, rtol=1e-3, atol=2e-4)     # world = 1: home IS the shard
    assert store.read("global/global_step").item() == 2 * nsteps
    assert st.step[:nsteps].cpu().tolist() == list(range(nsteps, 2 * nsteps))
    sg.close(); store.close()


def test_scratch_poisoning_changes_nothing(monkeypatch):
    """Race self-check of the fused step kernel (compute-sanitizer is closed on this pool): with DTFE_MLP2_POISON=1
    every scratch region - split-K partials, H, dA, dZ, row losses - is overwritten with NaN the moment its last
    legitimate reader is done.  A read that is not ordered after its producer (or before the next one) would pick
    up a NaN; results must stay bit-identical to the unpoisoned run, asynchronous and synchronous."""
    from distributedtensorflowexample_b200.parallel.sync import SyncGroup
    from distributedtensorflowexample_b200.train.gpu_step import FusedMlp2Stepper
    spec = reference_mlp_spec()
    x, y = _data(3000)
    X = torch.tensor(x, device="cuda"); Y = torch.tensor(y, device="cuda")
    idx = torch.tensor(np.random.default_rng(2).integers(0, 3000, size=40 * 100).astype(np.int32), device="cuda")
    for sync in (False, True):
        outs = []
        for poison in ("0", "1"):
            monkeypatch.setenv("DTFE_MLP2_POISON", poison)
            model, store, names = _store(spec, num_ps=2, seed=17)
            if sync:
                sg = SyncGroup(store, names, 0, 1, timeout_s=5.0, mode="allreduce_ll"); sg.refresh_replicas()
                st = FusedMlp2Stepper(spec, 100, sg.home_params(), sg.home("global/global_step"), 0.05, "cuda:0",
                                      max_steps_per_launch=20, remote=False,
                                      grad_buffers=sg.grad_views({n: tuple(store.tensor(n).shape) for n in names}))
                st.sync_exchange = sg.exchange_struct(0.05)
            else:
                st = FusedMlp2Stepper(spec, 100, [store.tensor(n) for n in names], store.tensor("global/global_step"),
                                      0.05, "cuda:0", max_steps_per_launch=20)
            losses = []
            for k in range(2):
                st.run(X, Y, nsteps=20, batch_idx=idx[k * 2000:(k + 1) * 2000])
                losses += st.loss[:20].cpu().tolist()
            torch.cuda.synchronize()
            outs.append((losses, [store.read(n) for n in names]))
            store.close()
        assert np.isfinite(outs[1][0]).all()
        assert outs[0][0] == outs[1][0]
        for a, b in zip(outs[0][1], outs[1][1]):
            assert np.array_equal(a, b)


def test_tracer_device_spans_time_the_step_kernel():
    """utils.tracing on a GPU: a device span is a CUDA-event pair on the launching stream (plus an NVTX range
    for the host span); its duration must be the kernel's, not the host call's."""
    from distributedtensorflowexample_b200.train.gpu_step import FusedMlp2Stepper
    from distributedtensorflowexample_b200.utils import tracing
    spec = reference_mlp_spec()
    model, store, names = _store(spec)
    x, y = _data()
    X = torch.tensor(x, device="cuda"); Y = torch.tensor(y, device="cuda")
    nsteps = 50
    st = FusedMlp2Stepper(spec, 100, [store.tensor(n) for n in names], store.tensor("global/global_step"), 0.01,
                          "cuda:0", max_steps_per_launch=nsteps)
    idx = torch.randint(0, x.shape[0], (nsteps * 100,), device="cuda", dtype=torch.int32)
    st.run(X, Y, nsteps=nsteps, batch_idx=idx)            # warm-up
    torch.cuda.synchronize()
    tr = tracing.Tracer(True, process_name="gpu-test")
    with tr.span("launch", "step", steps=nsteps), tr.device_span("step_kernel", steps=nsteps):
        st.run(X, Y, nsteps=nsteps, batch_idx=idx)
    evs = tr.events()
    dev = [e for e in evs if e["cat"] == "device"]
    host = [e for e in evs if e["name"] == "launch"]
    assert len(dev) == 1 and len(host) == 1
    per_step = dev[0]["dur"] / nsteps
    assert 5.0 < per_step < 200.0, per_step              # microseconds per training step, measured on the device
    assert dev[0]["args"] == {"steps": nsteps}
    store.close()
