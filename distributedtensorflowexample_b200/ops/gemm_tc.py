"""tcgen05 GEMM (``csrc/cuda/gemm_tc.cu``) - python face.

``gemm_tc(A, B, ...)`` exposes the raw operand-major interface; ``dense_fwd_tc`` /
``dense_bwd_dx_tc`` / ``dense_bwd_dw_push_tc`` are the three uses the MLP makes of it with the
reference's ``[in, out]`` kernel layout (reference worker.py:50-53, 72-79)."""
from __future__ import annotations

import ctypes
from ctypes import c_float, c_int, c_long, c_void_p
from typing import Optional

import torch

from . import cuda_ops
from ._lib import check

EPI_STORE, EPI_BIAS_ACT, EPI_ACTGRAD, EPI_PUSH = 0, 1, 2, 3


class _GemmTc(ctypes.Structure):
    _fields_ = [
        ("M", c_int), ("N", c_int), ("K", c_int),
        ("A", c_void_p), ("lda", c_long), ("a_mn_major", c_int),
        ("B", c_void_p), ("ldb", c_long), ("b_mn_major", c_int),
        ("D", c_void_p), ("ldd", c_long), ("out_bf16", c_int),
        ("epilogue", c_int),
        ("bias", c_void_p), ("bias_bf16", c_int), ("act", c_int),
        ("H", c_void_p), ("ldh", c_long),
        ("alpha", c_float), ("G", c_void_p), ("ldg", c_long),
        ("tile_n", c_int),
        ("nparts", c_int), ("rows_per_part", c_int),
        ("D_parts", c_void_p * 8),
        ("tma_push", c_int),
    ]


cuda_ops.register_sigs({"dtfe_gemm_tc_bf16": ([c_void_p, c_void_p], c_int),
                        "dtfe_gemm_tc_tf32": ([c_void_p, c_void_p], c_int)})


def _bf16(t: torch.Tensor, name: str, dtype=torch.bfloat16) -> torch.Tensor:
    if t.dtype != dtype or not t.is_cuda or t.dim() != 2 or t.stride(1) != 1:
        raise TypeError(f"{name}: expected a 2-D CUDA {dtype} tensor with contiguous rows, got {t.dtype} {tuple(t.shape)}")
    return t


def gemm_tc(A: torch.Tensor, B: torch.Tensor, M: int, N: int, K: int, a_mn_major: bool, b_mn_major: bool,
            D: Optional[torch.Tensor], epilogue: int = EPI_STORE, bias: Optional[torch.Tensor] = None, act: int = 0,
            H: Optional[torch.Tensor] = None, alpha: float = 0.0, G: Optional[torch.Tensor] = None,
            tile_n: int = 0, D_parts=None, tma_push: bool = False) -> None:
    """bf16 operands run as kind::f16, float32 operands as kind::tf32 (both accumulate in fp32)."""
    dt = A.dtype
    if dt not in (torch.bfloat16, torch.float32):
        raise TypeError("gemm_tc operands must be bfloat16 or float32")
    _bf16(A, "A", dt); _bf16(B, "B", dt)
    p = _GemmTc()
    p.M, p.N, p.K = M, N, K
    p.A, p.lda, p.a_mn_major = A.data_ptr(), A.stride(0), int(a_mn_major)
    p.B, p.ldb, p.b_mn_major = B.data_ptr(), B.stride(0), int(b_mn_major)
    if D is not None:
        if D.dtype not in (torch.float32, torch.bfloat16):
            raise TypeError("D must be float32 or bfloat16")
        p.D, p.ldd, p.out_bf16 = D.data_ptr(), D.stride(0), int(D.dtype == torch.bfloat16)
    p.epilogue = epilogue
    if bias is not None:
        p.bias, p.bias_bf16 = bias.data_ptr(), int(bias.dtype == torch.bfloat16)
    p.act = act
    if H is not None:
        _bf16(H, "H", dt)
        p.H, p.ldh = H.data_ptr(), H.stride(0)
    p.alpha = alpha
    if G is not None:
        p.G, p.ldg = G.data_ptr(), G.stride(0)
    p.tile_n = tile_n
    p.tma_push = int(tma_push)
    if D_parts is not None:
        p.nparts, p.rows_per_part = len(D_parts), D_parts[0].shape[0]
        p.ldd, p.out_bf16 = D_parts[0].stride(0), int(D_parts[0].dtype == torch.bfloat16)
        for i, t in enumerate(D_parts):
            p.D_parts[i] = t.data_ptr()
    fn = cuda_ops.lib().dtfe_gemm_tc_bf16 if dt == torch.bfloat16 else cuda_ops.lib().dtfe_gemm_tc_tf32
    check(fn(ctypes.byref(p), torch.cuda.current_stream().cuda_stream), "dtfe_gemm_tc")


def dense_fwd_tc(x, W, b, act: int, out, tile_n: int = 0):
    """out = act(x @ W + b); W [in,out] may be shard (peer) memory -> pull fused with the GEMM."""
    M, K = x.shape
    N = W.shape[1]
    gemm_tc(x, W, M, N, K, False, True, out, EPI_BIAS_ACT, bias=b, act=act, tile_n=tile_n)
    return out


def dense_bwd_dx_tc(dY, W, Hprev, out, tile_n: int = 0):
    """out = (dY @ W^T) * Hprev*(1-Hprev)."""
    M, N = dY.shape
    K = W.shape[0]
    gemm_tc(dY, W, M, K, N, False, False, out, EPI_ACTGRAD, H=Hprev, tile_n=tile_n)
    return out


def dense_bwd_dw_push_tc(X, dY, W, alpha: float, grad_out=None, tile_n: int = 0, tma_push: bool = False):
    """W += alpha * X^T @ dY, pushed from the GEMM epilogue with red.add (W may be peer memory, or a list of
    row blocks of a partitioned variable, each in a different shard)."""
    Mb, K = X.shape
    N = dY.shape[1]
    if isinstance(W, (list, tuple)):
        gemm_tc(X, dY, K, N, Mb, True, True, None, EPI_PUSH, alpha=alpha, G=grad_out, tile_n=tile_n, D_parts=list(W),
                tma_push=tma_push)
    else:
        gemm_tc(X, dY, K, N, Mb, True, True, W, EPI_PUSH, alpha=alpha, G=grad_out, tile_n=tile_n, tma_push=tma_push)


def smoke() -> None:
    g = torch.Generator(device="cuda").manual_seed(0)
    A = torch.randn(256, 192, device="cuda", generator=g).bfloat16()
    B = torch.randn(128, 192, device="cuda", generator=g).bfloat16()
    D = torch.empty(256, 128, device="cuda")
    gemm_tc(A, B, 256, 128, 192, False, False, D)
    torch.cuda.synchronize()
    ref = A.float() @ B.float().T
    err = (D - ref).abs().max().item()
    assert err < 1e-2 * ref.abs().max().item(), f"tcgen05 GEMM smoke mismatch: {err}"
