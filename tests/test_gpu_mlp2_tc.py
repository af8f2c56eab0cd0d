"""GPU: the tensor-core form of the fused MNIST step (csrc/cuda/mlp2_step_tc.cu) against the float64 oracle and
against the fp32 SIMT kernel.  tcgen05 kind::tf32 keeps 10 mantissa bits of X, W1 and dA, so the comparison is at
tf32 tolerance on the quantities those products feed (loss, parameter updates); bookkeeping is exact."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from distributedtensorflowexample_b200.models import oracle
from distributedtensorflowexample_b200.models.mlp import MLPSpec, reference_mlp_spec
from distributedtensorflowexample_b200.train.replica import build_replicated_model


def _store(spec, seed=0):
    from distributedtensorflowexample_b200.parallel.ps import PeerParamStore
    model = build_replicated_model(spec, num_ps=1)
    store = PeerParamStore(model.global_all, 1, torch.device("cuda:0"))
    store.allocate()
    store.initialize(np.random.default_rng(seed))
    return model, store, [v.name for v in model.global_trainable]


def _rel(a, b):
    return float(np.linalg.norm(a - b) / (np.linalg.norm(b) + 1e-30))


@pytest.mark.parametrize("dims,B", [((784, 100, 10), 100), ((52, 32, 7), 33), ((832, 96, 10), 104), ((100, 64, 10), 5),
                                    ((300, 20, 3), 64)])
def test_tc_step_matches_oracle_at_tf32_tolerance(dims, B):
    from distributedtensorflowexample_b200.data import synthetic_arrays
    from distributedtensorflowexample_b200.train.gpu_step import FusedMlp2Stepper
    spec = MLPSpec(dims)
    assert FusedMlp2Stepper.supports_tc(spec, B)
    model, store, names = _store(spec, seed=4)
    p0 = [store.read(n).astype(np.float64) for n in names]
    params_np = [p.copy() for p in p0]
    x, y = synthetic_arrays(1500, seed=5, num_classes=dims[2], dim=dims[0])
    y = y.astype(np.int32)
    X = torch.tensor(x, device="cuda"); Y = torch.tensor(y, device="cuda")
    lr, nsteps = 0.05, 6
    st = FusedMlp2Stepper(spec, B, [store.tensor(n) for n in names], store.tensor("global/global_step"), lr, "cuda:0",
                          max_steps_per_launch=nsteps, tc=True)
    assert st.tc
    idx = np.random.default_rng(1).integers(0, x.shape[0], size=2 * nsteps * B).astype(np.int32)
    idx_d = torch.tensor(idx, device="cuda")
    st.run(X, Y, nsteps=nsteps, batch_idx=idx_d[:nsteps * B])
    st.run(X, Y, nsteps=nsteps, batch_idx=idx_d[nsteps * B:])            # a second launch continues from the shard
    torch.cuda.synchronize()
    losses, accs = [], []
    for s in range(2 * nsteps):
        rows = idx[s * B:(s + 1) * B]
        l, a = oracle.sgd_step(params_np, x[rows].astype(np.float64), y[rows], lr)
        losses.append(l); accs.append(a)
    got_l = st.loss[:nsteps].cpu().numpy()
    np.testing.assert_allclose(got_l, losses[nsteps:], rtol=2e-2, atol=2e-2)
    assert np.abs(st.acc[:nsteps].cpu().numpy() - np.array(accs[nsteps:])).max() <= 2.0 / B + 1e-6   # an argmax may flip
    assert st.step[:nsteps].cpu().tolist() == list(range(nsteps, 2 * nsteps))
    assert store.read("global/global_step").item() == 2 * nsteps
    for n, p, q0 in zip(names, params_np, p0):
        got = store.read(n).astype(np.float64)
        assert _rel(got - q0, p - q0) < 3e-2, (n, _rel(got - q0, p - q0))   # the UPDATE is right to tf32 accuracy
        np.testing.assert_allclose(got, p, rtol=0, atol=2e-3 * max(1.0, float(np.abs(p).max())))
    store.close()


def test_tc_and_simt_kernels_agree_and_interleave():
    """Same shard, alternating launches of the two kernels: the step is the same step."""
    from distributedtensorflowexample_b200.data import synthetic_arrays
    from distributedtensorflowexample_b200.train.gpu_step import FusedMlp2Stepper
    spec = reference_mlp_spec()
    x, y = synthetic_arrays(2000, seed=9)
    y = y.astype(np.int32)
    X = torch.tensor(x, device="cuda"); Y = torch.tensor(y, device="cuda")
    idx = torch.tensor(np.random.default_rng(3).integers(0, 2000, size=40 * 100).astype(np.int32), device="cuda")
    finals = []
    for tc_pattern in ([False] * 4, [True] * 4, [True, False, True, False]):
        model, store, names = _store(spec, seed=8)
        args = ([store.tensor(n) for n in names], store.tensor("global/global_step"), 0.05, "cuda:0")
        simt = FusedMlp2Stepper(spec, 100, *args, max_steps_per_launch=10)
        tc = FusedMlp2Stepper(spec, 100, *args, max_steps_per_launch=10, tc=True)
        losses = []
        for i, use_tc in enumerate(tc_pattern):
            k = tc if use_tc else simt
            k.run(X, Y, nsteps=10, batch_idx=idx[i * 1000:(i + 1) * 1000])
            losses += k.loss[:10].cpu().tolist()
        torch.cuda.synchronize()
        assert store.read("global/global_step").item() == 40
        finals.append(([store.read(n).astype(np.float64) for n in names], losses))
        store.close()
    ref_p, ref_l = finals[0]
    assert ref_l[-1] < 0.5 * ref_l[0]                                  # it trains
    for p, l in finals[1:]:
        np.testing.assert_allclose(l, ref_l, rtol=3e-2, atol=3e-2)
        for a, b in zip(p, ref_p):
            np.testing.assert_allclose(a, b, rtol=0, atol=3e-3 * max(1.0, float(np.abs(b).max())))


def test_tc_rejects_what_it_does_not_cover():
    from distributedtensorflowexample_b200.ops import cuda_ops as ops
    assert ops.lib().dtfe_mlp2_step_tc_supported(100, 784, 100, 10) == 1
    assert ops.lib().dtfe_mlp2_step_tc_supported(128, 784, 100, 10) == 0       # batch beyond 104
    assert ops.lib().dtfe_mlp2_step_tc_supported(100, 784, 102, 10) == 0       # H not a multiple of 4
    assert ops.lib().dtfe_mlp2_step_tc_supported(100, 900, 100, 10) == 0       # more than 8 x 104 features
    assert ops.lib().dtfe_mlp2_step_tc_supported(104, 832, 128, 16) == 0       # small parameters do not fit next to the operands
