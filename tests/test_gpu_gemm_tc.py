"""GPU numerics of the tcgen05 GEMM against fp32 PyTorch on the same bf16-rounded operands."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _mk(shape, seed, scale=1.0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return (torch.randn(*shape, device="cuda", generator=g) * scale).bfloat16()


def _check(out, ref, K):
    tol = 2e-2 * (K ** 0.5) * 0.05 + 1e-2
    err = (out.float() - ref).abs().max().item()
    scale = ref.abs().max().item() + 1e-6
    assert err <= 1e-2 * scale + tol * 0 + 2e-2 * (1 if out.dtype == torch.bfloat16 else 0) * scale, (err, scale)


@pytest.mark.parametrize("tile_n", [128, 256, 512])
@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (256, 256, 256), (100, 100, 784), (300, 520, 200), (1024, 1024, 1024), (2048, 1536, 520)])
@pytest.mark.parametrize("a_mn,b_mn", [(0, 0), (0, 1), (1, 1), (1, 0)])
def test_gemm_tc_store_all_majors(M, N, K, a_mn, b_mn, tile_n):
    from distributedtensorflowexample_b200.ops import gemm_tc as G
    pad = lambda n: (n + 7) // 8 * 8
    A = _mk((K, pad(M)) if a_mn else (M, pad(K)), 1)
    B = _mk((K, pad(N)) if b_mn else (N, pad(K)), 2)
    Av = A[:, :M] if a_mn else A[:, :K]
    Bv = B[:, :N] if b_mn else B[:, :K]
    D = torch.full((M, N), float("nan"), device="cuda")
    G.gemm_tc(Av, Bv, M, N, K, bool(a_mn), bool(b_mn), D, tile_n=tile_n)
    torch.cuda.synchronize()
    Af = Av.float().T if a_mn else Av.float()
    Bf = Bv.float() if b_mn else Bv.float().T
    ref = Af @ Bf
    assert torch.isfinite(D).all()
    err = (D - ref).abs().max().item()
    assert err <= 2e-3 * ref.abs().max().item() + 1e-3, (err, ref.abs().max().item())


@pytest.mark.parametrize("B_,din,dout", [(100, 784, 104), (256, 512, 384), (4096, 784, 1024)])
def test_dense_layer_ops_tc(B_, din, dout):
    from distributedtensorflowexample_b200.ops import gemm_tc as G
    x = torch.rand(B_, din, device="cuda").bfloat16()
    W = _mk((din, dout), 3, 0.05)
    b = _mk((1, dout), 4)[0]
    out = torch.empty(B_, dout, device="cuda", dtype=torch.bfloat16)
    G.dense_fwd_tc(x, W, b, 1, out)
    ref = torch.sigmoid(x.float() @ W.float() + b.float())
    assert (out.float() - ref).abs().max().item() < 1e-2
    # dX with sigmoid-grad of the layer input
    dY = _mk((B_, dout), 5, 0.1)
    Hprev = torch.rand(B_, din, device="cuda").bfloat16()
    dX = torch.empty(B_, din, device="cuda", dtype=torch.bfloat16)
    G.dense_bwd_dx_tc(dY, W, Hprev, dX)
    ref_dx = (dY.float() @ W.float().T) * Hprev.float() * (1 - Hprev.float())
    assert (dX.float() - ref_dx).abs().max().item() < 2e-2 * ref_dx.abs().max().item() + 1e-3
    # dW push into an fp32 and a bf16 shard, plus the raw gradient
    ref_g = x.float().T @ dY.float()
    W32 = W.float().clone()
    Gout = torch.empty(din, dout, device="cuda")
    G.dense_bwd_dw_push_tc(x, dY, W32, -0.01, grad_out=Gout)
    assert (Gout - ref_g).abs().max().item() < 2e-3 * ref_g.abs().max().item() + 1e-3
    assert (W32 - (W.float() - 0.01 * ref_g)).abs().max().item() < 1e-3 * (1 + ref_g.abs().max().item() * 0.01)
    Wb = W.clone()
    G.dense_bwd_dw_push_tc(x, dY, Wb, -0.01)
    want = W.float() - 0.01 * ref_g
    assert (Wb.float() - want).abs().max().item() < 2e-2 * want.abs().max().item() + 1e-3


@pytest.mark.parametrize("B_,din,dout", [(4096, 1024, 2048), (4000, 1032, 2056)])
def test_forward_and_dx_epilogues_through_tma_store(B_, din, dout, monkeypatch):
    """DTFE_TMA_STORE=1: the 2-CTA kernel stages the bf16 output tile in shared memory and stores it with TMA
    (rows and columns beyond the matrix are clipped by the tensor map); same numbers as the per-thread stores."""
    from distributedtensorflowexample_b200.ops import gemm_tc as G
    x = torch.rand(B_, din, device="cuda").bfloat16()
    W = _mk((din, dout), 3, 0.05)
    b = _mk((1, dout), 4)[0]
    dY = _mk((B_, dout), 5, 0.1)
    Hprev = torch.rand(B_, din, device="cuda").bfloat16()
    outs = {}
    for mode in ("0", "1"):
        if mode == "1":
            monkeypatch.setenv("DTFE_TMA_STORE", "1")
        else:
            monkeypatch.delenv("DTFE_TMA_STORE", raising=False)
        out = torch.full((B_ + 8, dout), 7.0, device="cuda", dtype=torch.bfloat16)      # guard rows behind the matrix
        G.dense_fwd_tc(x, W, b, 1, out[:B_], tile_n=512)
        dX = torch.full((B_ + 8, din), 7.0, device="cuda", dtype=torch.bfloat16)
        G.dense_bwd_dx_tc(dY, W, Hprev, dX[:B_], tile_n=512)
        torch.cuda.synchronize()
        assert (out[B_:] == 7.0).all() and (dX[B_:] == 7.0).all()                       # nothing written out of bounds
        outs[mode] = (out[:B_].float(), dX[:B_].float())
    ref = torch.sigmoid(x.float() @ W.float() + b.float())
    assert (outs["1"][0] - ref).abs().max().item() < 1e-2
    assert torch.equal(outs["0"][0], outs["1"][0]) and torch.equal(outs["0"][1], outs["1"][1])


@pytest.mark.parametrize("parts", [0, 2, 4])
@pytest.mark.parametrize("M,N,K", [(1024, 512, 256), (2048, 1536, 1024), (1000, 520, 300)])
def test_push_epilogue_through_tma_reduce(M, N, K, parts):
    """dW push of the 2-CTA kernel: bf16 target, whole tiles added with cp.reduce.async.bulk.tensor; must
    agree with the per-thread red.add path and with fp32 math, also when the target is row-partitioned."""
    from distributedtensorflowexample_b200.ops import gemm_tc as G
    if parts and M % (parts * 128):
        pytest.skip("row blocks must be multiples of 128 rows for the TMA path (others fall back)")
    X = _mk((K, M), 11, 0.5)        # A is [K][M] (M contiguous), B is [K][N]
    dY = _mk((K, N), 12, 0.5)
    W0 = _mk((M, N), 13, 1.0)
    ref = W0.float() - 0.01 * (X.float().T @ dY.float())
    outs = []
    for no_tma in ("", "1"):
        use_tma = not no_tma
        W = W0.clone()
        if parts:
            rp = M // parts
            blocks = [W[i * rp:(i + 1) * rp] for i in range(parts)]
            G.gemm_tc(X, dY, M, N, K, True, True, None, G.EPI_PUSH, alpha=-0.01, tile_n=512, D_parts=blocks, tma_push=use_tma)
        else:
            G.gemm_tc(X, dY, M, N, K, True, True, W, G.EPI_PUSH, alpha=-0.01, tile_n=512, tma_push=use_tma)
        torch.cuda.synchronize()
        outs.append(W.float())
    tol = 2 ** -7 * ref.abs().max().item() + 1e-3
    assert (outs[0] - ref).abs().max().item() <= tol
    assert (outs[1] - ref).abs().max().item() <= tol


@pytest.mark.parametrize("tile_n", [128, 512])
@pytest.mark.parametrize("M,N,K", [(128, 128, 32), (100, 100, 784), (300, 520, 200), (1024, 768, 1024)])
@pytest.mark.parametrize("a_mn,b_mn", [(0, 0), (0, 1), (1, 1), (1, 0)])
def test_gemm_tc_tf32_all_majors(M, N, K, a_mn, b_mn, tile_n):
    """fp32 operands through kind::tf32 (10-bit mantissa inputs, fp32 accumulation)."""
    from distributedtensorflowexample_b200.ops import gemm_tc as G
    pad = lambda n: (n + 3) // 4 * 4
    g = torch.Generator(device="cuda").manual_seed(7)
    A = torch.randn((K, pad(M)) if a_mn else (M, pad(K)), device="cuda", generator=g)
    B = torch.randn((K, pad(N)) if b_mn else (N, pad(K)), device="cuda", generator=g)
    Av = A[:, :M] if a_mn else A[:, :K]
    Bv = B[:, :N] if b_mn else B[:, :K]
    D = torch.full((M, N), float("nan"), device="cuda")
    G.gemm_tc(Av, Bv, M, N, K, bool(a_mn), bool(b_mn), D, tile_n=tile_n)
    torch.cuda.synchronize()
    Af = Av.double().T if a_mn else Av.double()
    Bf = Bv.double() if b_mn else Bv.double().T
    ref = Af @ Bf
    assert torch.isfinite(D).all()
    err = (D.double() - ref).abs().max().item()
    # tf32 operands are TRUNCATED to 10 mantissa bits by the tensor core: compare with that model exactly,
    # and with fp64 loosely
    t32 = lambda x: (x.contiguous().view(torch.int32) & ~0x1fff).view(torch.float32).double()
    ref_t = (t32(Av).T if a_mn else t32(Av)) @ (t32(Bv) if b_mn else t32(Bv).T)
    assert (D.double() - ref_t).abs().max().item() <= 1e-3 * (1 + ref.abs().max().item() * 1e-2)
    assert err <= 5e-3 * (K ** 0.5) + 1e-3, (err, ref.abs().max().item())


def test_dense_layer_ops_tf32_on_reference_shapes():
    """The reference model's own shapes and dtype (fp32 [784,100] kernel, batch 100) on the tensor cores."""
    from distributedtensorflowexample_b200.ops import gemm_tc as G
    g = torch.Generator(device="cuda").manual_seed(3)
    x = torch.rand(100, 784, device="cuda", generator=g)
    W = torch.randn(784, 100, device="cuda", generator=g) * 0.05
    b = torch.randn(100, device="cuda", generator=g)
    out = torch.empty(100, 100, device="cuda")
    G.dense_fwd_tc(x, W, b, 1, out)
    ref = torch.sigmoid(x.double() @ W.double() + b.double())
    assert (out.double() - ref).abs().max().item() < 5e-3
    dY = torch.randn(100, 100, device="cuda", generator=g) * 0.1
    Wn = W.clone()
    G.dense_bwd_dw_push_tc(x, dY, Wn, -0.1)
    want = W.double() - 0.1 * (x.double().T @ dY.double())
    assert (Wn.double() - want).abs().max().item() < 5e-3
    Hprev = torch.rand(100, 784, device="cuda", generator=g)
    dX = torch.empty(100, 784, device="cuda")
    G.dense_bwd_dx_tc(dY, W, Hprev, dX)
    ref_dx = (dY.double() @ W.double().T) * Hprev.double() * (1 - Hprev.double())
    assert (dX.double() - ref_dx).abs().max().item() < 5e-3
