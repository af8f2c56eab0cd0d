"""The comparison baseline's arithmetic on the CPU (world = 1): its flat-buffer form, its per-variable form and the
numpy oracle must be the same SGD step - otherwise `vs_baseline` would compare against a different algorithm."""
import numpy as np
import torch

from baseline.nccl_ps import NcclPsWorker
from distributedtensorflowexample_b200.models import oracle


def test_flat_and_per_variable_baseline_match_the_oracle():
    dev = torch.device("cpu")
    dims, B, lr = (784, 100, 10), 100, 0.05
    g = torch.Generator().manual_seed(3)
    x = torch.rand(5 * B, dims[0], generator=g)
    y = torch.randint(0, 10, (5 * B,), generator=g, dtype=torch.int32)
    workers = [NcclPsWorker(dims, B, lr, 0, 1, dev, torch.float32, flat=f) for f in (True, False)]
    params_np = [p.detach().numpy().astype(np.float64).copy() for p in workers[0].params]
    for a, b in zip(workers[0].params, workers[1].params):
        assert torch.equal(a, b)                       # same seed, same init
    for s in range(5):
        xb, yb = x[s * B:(s + 1) * B], y[s * B:(s + 1) * B]
        for w in workers:
            w.step(xb, yb)
        loss, _ = oracle.sgd_step(params_np, xb.numpy().astype(np.float64), yb.numpy().astype(np.int64), lr)
        assert abs(float(workers[0].loss) - loss) < 1e-3 * max(1.0, loss)
    for a, b, ref in zip(workers[0].params, workers[1].params, params_np):
        np.testing.assert_allclose(a.numpy(), b.numpy(), rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(a.numpy(), ref, rtol=1e-3, atol=2e-4)
    assert int(workers[0].global_step) == 5 == int(workers[1].global_step)
    # the variables of the flat form are views of one buffer (one all-reduce moves all gradients)
    base = workers[0].flat_p.data_ptr()
    assert all(base <= p.data_ptr() < base + workers[0].flat_p.numel() * 4 for p in workers[0].params)
