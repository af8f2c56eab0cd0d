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


@pytest.mark.parametrize("M,K,C", [(300, 1024, 10), (64, 256, 7), (4096, 2048, 16), (33, 8, 1)])
def test_thin_layer_kernels_match_fp32(M, K, C):
    """The streaming kernels of an output layer of width <= 16 against plain fp32 torch on the same bf16 values."""
    from distributedtensorflowexample_b200.ops import cuda_ops as ops
    g = torch.Generator(device="cuda").manual_seed(M + K + C)
    pitch = 16 if C > 8 else 8
    act = torch.rand(M, K, device="cuda", generator=g).bfloat16()
    Wp = torch.zeros(K, pitch, device="cuda", dtype=torch.bfloat16)
    Wp[:, :C] = (torch.randn(K, C, device="cuda", generator=g) * 0.05).bfloat16()
    W = Wp[:, :C]
    bias = torch.randn(C, device="cuda", generator=g).bfloat16()
    dZp = torch.zeros(M, pitch, device="cuda", dtype=torch.bfloat16)
    dZp[:, :C] = (torch.randn(M, C, device="cuda", generator=g) / M).bfloat16()
    dZ = dZp[:, :C]
    assert ops.thin_ok(act, W)
    out = torch.empty(M, pitch, device="cuda")[:, :C]
    ops.thin_fwd(act, W, bias, out)
    ref = act.float() @ W.float() + bias.float()
    torch.testing.assert_close(out, ref, rtol=1e-4, atol=1e-3)
    dX = torch.empty(M, K, device="cuda", dtype=torch.bfloat16)
    ops.thin_dx(dZ, W, act, dX)
    ref = (dZ.float() @ W.float().t()) * act.float() * (1 - act.float())
    torch.testing.assert_close(dX.float(), ref, rtol=2e-2, atol=float(ref.abs().max()) * 2 ** -7)
    scratch = ops.thin_dw_scratch(M, K, "cuda")
    before = Wp.clone()
    for rep in range(2):          # twice: the kernel must leave its counters ready for the next launch
        ops.thin_dw_push(act, dZ, W, -0.5, scratch)
    torch.cuda.synchronize()
    upd = -0.5 * (act.float().t() @ dZ.float())
    got = (Wp.float() - before.float())[:, :C]
    assert (got - 2 * upd).abs().max() <= 2 ** -6 * before.float().abs().max() + 0.03 * upd.abs().max() + 1e-6
    assert torch.equal(Wp[:, C:], before[:, C:])          # pad columns untouched (adding zero)
