"""GPU: the bf16 tcgen05 step engine against the fp32 oracle run on the same bf16-rounded parameters."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from distributedtensorflowexample_b200.models import oracle
from distributedtensorflowexample_b200.models.mlp import MLPSpec
from distributedtensorflowexample_b200.train.replica import build_replicated_model


@pytest.mark.parametrize("partition", [0, 1])
@pytest.mark.parametrize("dims,B", [((784, 256, 128, 10), 128), ((784, 512, 10), 200)])
def test_tc_step_matches_oracle(dims, B, partition):
    from distributedtensorflowexample_b200.parallel.ps import PeerParamStore
    from distributedtensorflowexample_b200.train.gpu_step import TcStepper
    from distributedtensorflowexample_b200.data import synthetic_arrays
    spec = MLPSpec(dims, param_dtype="bfloat16")
    model = build_replicated_model(spec, num_ps=2)
    store = PeerParamStore(model.global_all, 2, torch.device("cuda:0"), partition_min_bytes=(64 << 10) if partition else 0)
    store.allocate()
    rng = np.random.default_rng(0)
    names = [v.name for v in model.global_trainable]
    for v in model.global_trainable:      # small weights so sigmoids are not saturated and bf16 error stays visible
        store.assign(v.name, (rng.normal(size=v.shape) * 0.05).astype(np.float32))
    W_last = store.tensor(names[-2])
    assert W_last.shape == (dims[-2], 10) and W_last.stride(0) == 16          # padded pitch inside the shard
    assert store.is_partitioned(names[0]) == bool(partition)                   # W1 (>= 64 KB) is split by rows
    if partition:
        blocks = store.parts(names[0])
        assert len(blocks) == 2 and blocks[0].shape == (392, dims[1]) and store.shard_of(names[0]) == -1
    params_np = [store.read(n).astype(np.float64) for n in names]              # bf16-rounded values
    x, y = synthetic_arrays(B, seed=3)
    xb = torch.tensor(x, device="cuda").bfloat16()
    X = xb
    Y = torch.tensor(y.astype(np.int32), device="cuda")
    lr = 0.5
    st = TcStepper(spec, B, [store.param_handle(n) for n in names], store.tensor("global/global_step"), lr, "cuda:0")
    st.run(X, Y)
    torch.cuda.synchronize()
    loss, acc, grads = oracle.gradients(params_np, xb.float().cpu().numpy().astype(np.float64), y.astype(np.int64))
    assert abs(st.loss.item() - loss) < 2e-2 * max(1.0, loss)
    assert abs(st.acc.item() - acc) <= 2.0 / B
    assert st.step.item() == 0 and store.read("global/global_step").item() == 1
    for n, p, g in zip(names, params_np, grads):
        want = p - lr * g
        got = store.read(n).astype(np.float64)
        upd_err = np.abs((got - p) - (-lr * g)).max()
        # bf16 storage rounds the updated weight (<= 2^-8 relative) and the GEMM inputs are bf16
        assert upd_err <= 2 ** -7 * np.abs(want).max() + 0.05 * np.abs(lr * g).max() + 1e-4, (n, upd_err)
    store.close()
