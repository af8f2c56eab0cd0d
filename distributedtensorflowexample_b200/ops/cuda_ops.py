"""Python face of ``libdtfe_cuda.so``: torch tensors in, raw pointers + the current CUDA
stream out.  No op here has an eager-PyTorch fallback - if the library is missing the
import of the symbol table raises.

Every wrapper cites the reference op it replaces (file:line in /root/reference).
"""
from __future__ import annotations

import ctypes
from ctypes import c_int, c_long, c_size_t, c_float, c_void_p, c_uint, POINTER
from typing import Optional, Sequence

import torch

from ._lib import cuda_lib, check

_P = c_void_p

_SIGS = {
    "dtfe_abi_version": ([], c_int),
    "dtfe_struct_sizes": ([POINTER(c_size_t), c_int], c_int),
    "dtfe_device_count": ([POINTER(c_int)], c_int),
    "dtfe_device_info": ([c_int, POINTER(c_int)], c_int),
    "dtfe_shard_alloc": ([c_size_t, POINTER(_P)], c_int),
    "dtfe_shard_free": ([_P], c_int),
    "dtfe_ipc_export": ([_P, _P], c_int),
    "dtfe_ipc_open": ([_P, POINTER(_P)], c_int),
    "dtfe_ipc_close": ([_P], c_int),
    "dtfe_enable_peer": ([c_int, c_int], c_int),
    "dtfe_l2_flush": ([_P, c_size_t, _P], c_int),
    "dtfe_global_step_add": ([_P, c_int, _P, _P], c_int),
    "dtfe_dense_fwd_f32": ([_P, c_long, _P, c_long, _P, _P, c_long, c_int, c_int, c_int, c_int, c_int, _P], c_int),
    "dtfe_dense_bwd_dx_f32": ([_P, c_long, _P, c_long, _P, c_long, _P, c_long, c_int, c_int, c_int, c_int, c_int, _P], c_int),
    "dtfe_dense_bwd_dw_push_f32": ([_P, c_long, _P, c_long, _P, c_long, _P, c_long, c_int, c_int, c_int, c_float, c_int, _P], c_int),
    "dtfe_bias_grad_push_f32": ([_P, c_long, c_int, c_int, _P, _P, c_float, c_int, _P], c_int),
    "dtfe_softmax_xent_f32": ([_P, c_long, _P, _P, c_long, c_int, c_int, c_float, _P, c_long, _P, _P, _P], c_int),
    "dtfe_softmax_xent_f32_bf16grad": ([_P, c_long, _P, c_int, c_int, c_float, _P, c_long, _P, c_long, _P, _P, _P], c_int),
    "dtfe_bias_grad_push_bf16": ([_P, c_long, c_int, c_int, _P, c_int, _P, c_float, _P], c_int),
    "dtfe_mean_loss_acc": ([_P, _P, c_int, _P, _P, _P, _P], c_int),
    "dtfe_push_sgd_f32": ([_P, _P, c_size_t, c_float, c_int, _P], c_int),
    "dtfe_push_sgd_bf16": ([_P, _P, c_size_t, c_float, _P], c_int),
    "dtfe_pull_f32": ([_P, _P, c_size_t, _P], c_int),
    "dtfe_memcpy_async": ([_P, _P, c_size_t, _P], c_int),
    "dtfe_adam_apply_f32": ([_P, _P, _P, _P, c_size_t, c_float, c_float, c_float, c_float, c_float, _P], c_int),
    "dtfe_sync_rs_apply_f32": ([POINTER(_P), POINTER(_P), c_int, c_int, c_int, POINTER(_P), POINTER(c_size_t),
                                POINTER(c_size_t), c_float, c_uint, _P, _P, _P, ctypes.c_double, _P], c_int),
    "dtfe_sync_max_ranks": ([], c_int),
    "dtfe_gather_rows_f32": ([_P, c_long, _P, _P, _P, c_int, c_int, _P, _P, _P], c_int),
    "dtfe_thin_fwd_bf16": ([_P, c_long, _P, c_long, _P, _P, c_long, c_int, c_int, c_int, _P], c_int),
    "dtfe_thin_dx_bf16": ([_P, c_long, _P, c_long, _P, c_long, _P, c_long, c_int, c_int, c_int, _P], c_int),
    "dtfe_thin_dw_scratch_words": ([c_int, c_int], c_size_t),
    "dtfe_thin_dw_push_bf16": ([_P, c_long, _P, c_long, _P, c_long, c_float, c_int, c_int, c_int, _P, _P], c_int),
    "dtfe_mlp2_step_grid": ([c_int, c_int], c_int),
    "dtfe_mlp2_step_ctas": ([c_int, c_int, c_int, c_int], c_int),
    "dtfe_mlp2_step_workspace_words": ([c_int, c_int, c_int, c_int, c_int], c_size_t),
    "dtfe_mlp2_step": ([_P, _P], c_int),
    "dtfe_mlp2_step_tc": ([_P, _P], c_int),
    "dtfe_mlp2_step_tc_supported": ([c_int, c_int, c_int, c_int], c_int),
    "dtfe_hostfeed_create": ([c_int, c_int, c_int, POINTER(_P)], c_int),
    "dtfe_hostfeed_destroy": ([_P], c_int),
    "dtfe_mlp2_train_host": ([_P, _P, POINTER(_P), POINTER(_P), c_int, _P, _P], c_int),
}

_lib = None


def lib():
    global _lib
    if _lib is None:
        L = cuda_lib()
        for name, (argtypes, restype) in _SIGS.items():
            fn = getattr(L, name)
            fn.argtypes = argtypes
            fn.restype = restype
        _lib = L
    return _lib


def register_sigs(sigs: dict) -> None:
    """Other op modules (gemm_tc, ...) add their entry points to the same table."""
    global _lib
    _SIGS.update(sigs)
    _lib = None


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _f32c(t: torch.Tensor, name: str) -> torch.Tensor:
    if t.dtype != torch.float32 or not t.is_cuda:
        raise TypeError(f"{name}: expected a CUDA float32 tensor, got {t.dtype} on {t.device}")
    if t.dim() == 2 and t.stride(1) != 1:
        raise ValueError(f"{name}: inner dimension must be contiguous")
    return t


# --------------------------------------------------------------------------------------
# raw-pointer tensors: shards are cudaMalloc'd by the library (so they can be exported over
# CUDA IPC) and viewed as torch tensors without a copy.
# --------------------------------------------------------------------------------------
class _RawCudaBuffer:
    def __init__(self, ptr: int, nbytes: int, owner=None):
        self.ptr, self.nbytes, self.owner = ptr, nbytes, owner
        self.__cuda_array_interface__ = {
            "shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3, "strides": None,
        }


def tensor_from_ptr(ptr: int, shape: Sequence[int], dtype: torch.dtype, device: torch.device, owner=None) -> torch.Tensor:
    n = 1
    for s in shape:
        n *= int(s)
    nbytes = n * torch.empty((), dtype=dtype).element_size()
    raw = torch.as_tensor(_RawCudaBuffer(ptr, max(nbytes, 1), owner), device=device)
    return raw[:nbytes].view(dtype).view(*shape)


def shard_alloc(nbytes: int) -> int:
    out = _P()
    check(lib().dtfe_shard_alloc(nbytes, ctypes.byref(out)), "dtfe_shard_alloc")
    return out.value


def shard_free(ptr: int) -> None:
    check(lib().dtfe_shard_free(ptr), "dtfe_shard_free")


def ipc_export(ptr: int) -> bytes:
    buf = ctypes.create_string_buffer(64)
    check(lib().dtfe_ipc_export(ptr, buf), "dtfe_ipc_export")
    return buf.raw


def ipc_open(handle: bytes) -> int:
    out = _P()
    buf = ctypes.create_string_buffer(handle, 64)
    check(lib().dtfe_ipc_open(buf, ctypes.byref(out)), "dtfe_ipc_open")
    return out.value


def ipc_close(ptr: int) -> None:
    check(lib().dtfe_ipc_close(ptr), "dtfe_ipc_close")


def enable_peer(dev: int, peer: int) -> None:
    check(lib().dtfe_enable_peer(dev, peer), "dtfe_enable_peer")


def device_info(dev: int = 0) -> dict:
    out = (c_int * 6)()
    check(lib().dtfe_device_info(dev, out), "dtfe_device_info")
    return dict(major=out[0], minor=out[1], sm_count=out[2], smem_optin=out[3], l2_bytes=out[4], coop=out[5])


def l2_flush(buf: torch.Tensor) -> None:
    check(lib().dtfe_l2_flush(buf.data_ptr(), buf.numel() * buf.element_size(), _stream()), "dtfe_l2_flush")


def global_step_add(gs: torch.Tensor, inc: int = 1, observed: Optional[torch.Tensor] = None) -> None:
    """reference worker.py:32,141 ``global_step.assign_add(1)`` as a system-scope atomic."""
    check(lib().dtfe_global_step_add(gs.data_ptr(), inc, _ptr(observed), _stream()), "dtfe_global_step_add")


# --------------------------------------------------------------------------------------
# dense layers (fp32 SIMT path)
# --------------------------------------------------------------------------------------
ACT_NONE, ACT_SIGMOID = 0, 1


def dense_fwd(x, W, b, act: int, out=None, shard_reads: bool = True):
    """``act(x @ W + b)`` - reference worker.py:50-53 (tf.layers.dense, kernel layout [in,out])."""
    _f32c(x, "x"); _f32c(W, "W")
    M, K = x.shape
    K2, N = W.shape
    assert K == K2, (x.shape, W.shape)
    if out is None:
        out = torch.empty(M, N, device=x.device, dtype=torch.float32)
    check(lib().dtfe_dense_fwd_f32(x.data_ptr(), x.stride(0), W.data_ptr(), W.stride(0), _ptr(b), out.data_ptr(),
                                   out.stride(0), M, N, K, act, int(shard_reads), _stream()), "dtfe_dense_fwd_f32")
    return out


def dense_bwd_dx(dY, W, Hprev, act: int, out=None, shard_reads: bool = True):
    """``(dY @ W^T) * act'(Hprev)`` - the MatMul-grad + SigmoidGrad pair of tf.gradients (worker.py:72-73)."""
    _f32c(dY, "dY"); _f32c(W, "W")
    M, N = dY.shape
    K, N2 = W.shape
    assert N == N2
    if out is None:
        out = torch.empty(M, K, device=dY.device, dtype=torch.float32)
    check(lib().dtfe_dense_bwd_dx_f32(dY.data_ptr(), dY.stride(0), W.data_ptr(), W.stride(0), _ptr(Hprev),
                                      Hprev.stride(0) if Hprev is not None else 0, out.data_ptr(), out.stride(0),
                                      M, N, K, act if Hprev is not None else 0, int(shard_reads), _stream()),
          "dtfe_dense_bwd_dx_f32")
    return out


def dense_bwd_dw_push(X, dY, W, alpha: float, atomic: bool = True, grad_out=None):
    """``W += alpha * X^T @ dY`` with the push fused into the GEMM epilogue (worker.py:75-79).
    ``W`` may be None (only ``grad_out`` is written: sync mode's local gradient buffer)."""
    _f32c(X, "X"); _f32c(dY, "dY")
    M, K = X.shape
    M2, N = dY.shape
    assert M == M2
    check(lib().dtfe_dense_bwd_dw_push_f32(X.data_ptr(), X.stride(0), dY.data_ptr(), dY.stride(0), _ptr(W),
                                           W.stride(0) if W is not None else 0, _ptr(grad_out),
                                           grad_out.stride(0) if grad_out is not None else 0, M, N, K, alpha,
                                           int(atomic), _stream()), "dtfe_dense_bwd_dw_push_f32")


def bias_grad_push(dY, b, alpha: float, atomic: bool = True, grad_out=None):
    """``b += alpha * dY.sum(0)`` (TF1 BiasAddGrad + ApplyGradientDescent)."""
    M, N = dY.shape
    check(lib().dtfe_bias_grad_push_f32(dY.data_ptr(), dY.stride(0), M, N, _ptr(b), _ptr(grad_out), alpha,
                                        int(atomic), _stream()), "dtfe_bias_grad_push_f32")


def softmax_xent(logits, labels, scale: float, dZ=None, rowloss=None, rowcorrect=None):
    """Fused softmax cross-entropy fwd+bwd and per-row correctness (worker.py:63-66, 88-90).
    ``labels``: int32 class ids [M] or a dense float32 [M,C] (one-hot / soft) matrix."""
    M, C = logits.shape
    ids = labels if labels.dtype == torch.int32 else None
    dense = labels if labels.dtype == torch.float32 else None
    if ids is None and dense is None:
        raise TypeError("labels must be int32 ids or float32 dense")
    check(lib().dtfe_softmax_xent_f32(logits.data_ptr(), logits.stride(0), _ptr(ids), _ptr(dense),
                                      dense.stride(0) if dense is not None else 0, M, C, scale, _ptr(dZ),
                                      dZ.stride(0) if dZ is not None else 0, _ptr(rowloss), _ptr(rowcorrect),
                                      _stream()), "dtfe_softmax_xent_f32")


def softmax_xent_bf16grad(logits, label_ids, scale: float, dZ, dZ_bf16, rowloss=None, rowcorrect=None):
    """softmax_xent that also emits the bf16 dZ operand of the tcgen05 backward GEMMs."""
    M, C = logits.shape
    check(lib().dtfe_softmax_xent_f32_bf16grad(logits.data_ptr(), logits.stride(0), label_ids.data_ptr(), M, C, scale,
                                               _ptr(dZ), dZ.stride(0) if dZ is not None else 0, dZ_bf16.data_ptr(),
                                               dZ_bf16.stride(0), _ptr(rowloss), _ptr(rowcorrect), _stream()),
          "dtfe_softmax_xent_f32_bf16grad")


def bias_grad_push_bf16(dY, b, alpha: float, grad_out=None):
    M, N = dY.shape
    check(lib().dtfe_bias_grad_push_bf16(dY.data_ptr(), dY.stride(0), M, N, _ptr(b),
                                         int(b is not None and b.dtype == torch.bfloat16), _ptr(grad_out), alpha,
                                         _stream()), "dtfe_bias_grad_push_bf16")


# --------------------------------------------------------------------------------------
# thin bf16 layer (output width <= 16): streaming kernels instead of 128-wide tensor-core tiles
# --------------------------------------------------------------------------------------
def thin_ok(act: torch.Tensor, W: torch.Tensor) -> bool:
    return (isinstance(W, torch.Tensor) and W.dim() == 2 and W.shape[1] <= 16 and W.dtype == torch.bfloat16
            and act.dtype == torch.bfloat16 and act.shape[1] % 8 == 0 and act.stride(0) % 8 == 0 and W.stride(0) % 8 == 0
            and (W.shape[1] <= 8 or W.stride(0) >= 16) and act.data_ptr() % 16 == 0 and W.data_ptr() % 16 == 0)


def thin_fwd(act, W, bias, out):
    """out[M,C] (fp32) = act[M,K] @ W[K,C] + bias - reference worker.py:53 for an output layer of width <= 16."""
    M, K = act.shape
    check(lib().dtfe_thin_fwd_bf16(act.data_ptr(), act.stride(0), W.data_ptr(), W.stride(0), _ptr(bias), out.data_ptr(),
                                   out.stride(0), M, K, W.shape[1], _stream()), "dtfe_thin_fwd_bf16")
    return out


def thin_dx(dZ, W, Hact, out):
    """out[M,K] (bf16) = (dZ[M,C] @ W[K,C]^T) * Hact*(1-Hact)."""
    M, K = Hact.shape
    check(lib().dtfe_thin_dx_bf16(dZ.data_ptr(), dZ.stride(0), W.data_ptr(), W.stride(0), Hact.data_ptr(), Hact.stride(0),
                                  out.data_ptr(), out.stride(0), M, K, W.shape[1], _stream()), "dtfe_thin_dx_bf16")
    return out


def thin_dw_scratch(M: int, K: int, device) -> torch.Tensor:
    return torch.zeros(lib().dtfe_thin_dw_scratch_words(M, K), dtype=torch.float32, device=device)


def thin_dw_push(act, dZ, Wshard, alpha: float, scratch):
    """Wshard[K,C] += alpha * act[M,K]^T @ dZ[M,C] (one rounded red.add per element into the bf16 shard)."""
    M, K = act.shape
    check(lib().dtfe_thin_dw_push_bf16(act.data_ptr(), act.stride(0), dZ.data_ptr(), dZ.stride(0), Wshard.data_ptr(),
                                       Wshard.stride(0), alpha, M, K, Wshard.shape[1], scratch.data_ptr(), _stream()),
          "dtfe_thin_dw_push_bf16")


def mean_loss_acc(rowloss, rowcorrect, M: int, out_loss=None, out_acc=None, out_count=None):
    check(lib().dtfe_mean_loss_acc(_ptr(rowloss), _ptr(rowcorrect), M, _ptr(out_loss), _ptr(out_acc),
                                   _ptr(out_count), _stream()), "dtfe_mean_loss_acc")


# --------------------------------------------------------------------------------------
# flat push / pull / optimizer kernels
# --------------------------------------------------------------------------------------
def push_sgd(shard: torch.Tensor, grad: torch.Tensor, alpha: float, atomic: bool = True):
    n = grad.numel()
    if shard.dtype == torch.float32:
        check(lib().dtfe_push_sgd_f32(shard.data_ptr(), grad.data_ptr(), n, alpha, int(atomic), _stream()),
              "dtfe_push_sgd_f32")
    elif shard.dtype == torch.bfloat16:
        check(lib().dtfe_push_sgd_bf16(shard.data_ptr(), grad.data_ptr(), n, alpha, _stream()), "dtfe_push_sgd_bf16")
    else:
        raise TypeError(shard.dtype)


def pull(dst: torch.Tensor, shard: torch.Tensor):
    check(lib().dtfe_pull_f32(dst.data_ptr(), shard.data_ptr(), dst.numel(), _stream()), "dtfe_pull_f32")


def pull_ce(dst: torch.Tensor, shard: torch.Tensor, nbytes: int) -> None:
    """Bulk pull on a copy engine (the current stream's): ``nbytes`` from ``shard`` (possibly peer HBM) to ``dst``."""
    check(lib().dtfe_memcpy_async(dst.data_ptr(), shard.data_ptr(), nbytes, _stream()), "dtfe_memcpy_async")


def adam_apply(param, m, v, slot, lr_t: float, beta1: float, beta2: float, eps: float, grad_scale: float = 1.0):
    check(lib().dtfe_adam_apply_f32(param.data_ptr(), m.data_ptr(), v.data_ptr(), slot.data_ptr(), param.numel(),
                                    lr_t, beta1, beta2, eps, grad_scale, _stream()), "dtfe_adam_apply_f32")


def gather_rows(X, labels, idx, cursor, Xout, Yout):
    """On-device ``next_batch``: rows ``idx[cursor*B : (cursor+1)*B]`` of the resident dataset (cursor on device)."""
    B, D = Xout.shape
    check(lib().dtfe_gather_rows_f32(X.data_ptr(), X.stride(0), labels.data_ptr(), idx.data_ptr(), cursor.data_ptr(),
                                     B, D, Xout.data_ptr(), Yout.data_ptr(), _stream()), "dtfe_gather_rows_f32")


def sync_rs_apply(grad_ptrs: Sequence[int], flag_ptrs: Sequence[int], rank: int, world: int,
                  segments: Sequence, alpha: float, epoch: int, local_sync: torch.Tensor,
                  global_step: Optional[torch.Tensor], abort_flag_ptr: Optional[int] = None, timeout_s: float = 20.0):
    """Synchronous SGD: reduce-scatter of all workers' gradients fused with the apply.
    ``segments``: (param_tensor, begin, end) for every variable this rank owns."""
    G = (_P * world)(*grad_ptrs)
    F = (_P * world)(*flag_ptrs)
    n = len(segments)
    P = (_P * max(n, 1))(*[t.data_ptr() for t, _, _ in segments])
    Bg = (c_size_t * max(n, 1))(*[b for _, b, _ in segments])
    En = (c_size_t * max(n, 1))(*[e for _, _, e in segments])
    check(lib().dtfe_sync_rs_apply_f32(G, F, rank, world, n, P, Bg, En, alpha, epoch, local_sync.data_ptr(),
                                       _ptr(global_step), abort_flag_ptr, timeout_s, _stream()),
          "dtfe_sync_rs_apply_f32")


# --------------------------------------------------------------------------------------
# fused 2-layer-MLP step (the reference model's whole step in one cooperative kernel)
# --------------------------------------------------------------------------------------
class SyncExchange(ctypes.Structure):
    """ctypes mirror of ``dtfe_sync_exchange`` (include/dtfe_b200.h)."""
    _fields_ = [
        ("grads", _P * 16), ("flags", _P * 16),
        ("rank", c_int), ("world", c_int),
        ("nseg", c_int),
        ("seg_param", _P * 8), ("seg_begin", c_size_t * 8), ("seg_end", c_size_t * 8),
        ("alpha", c_float),
        ("epoch", _P), ("status", _P), ("abort_flag", _P),
        ("timeout_ns", ctypes.c_ulonglong),
        ("global_step", _P),
        ("mode", c_int), ("grad_stride", c_size_t), ("cta_flags", _P * 16), ("ll_recv", _P * 16),
    ]


class _Mlp2Launch(ctypes.Structure):
    _fields_ = [
        ("B", c_int), ("D", c_int), ("H", c_int), ("C", c_int), ("ks", c_int),
        ("W1", _P), ("b1", _P), ("W2", _P), ("b2", _P),
        ("global_step", _P),
        ("X", _P), ("ldx", c_long),
        ("labels", _P),
        ("batch_idx", _P),
        ("rows_per_step", c_long),
        ("workspace", _P),
        ("loss_out", _P), ("acc_out", _P), ("step_out", _P),
        ("lr", c_float), ("nsteps", c_int), ("atomic", c_int), ("train", c_int), ("remote", c_int),
        ("dbg", _P),
        ("gW1", _P), ("gb1", _P), ("gW2", _P), ("gb2", _P),
        ("cursor", _P),
        ("sync_exchange", _P),
    ]


class Mlp2Step:
    """Owns the scratch workspace and per-step output rings of the fused step kernel."""

    def __init__(self, B: int, D: int, H: int, C: int, device, max_steps: int = 1, ks: int = 8):
        self.B, self.D, self.H, self.C, self.ks = B, D, H, C, ks
        self.device = torch.device(device)
        words = lib().dtfe_mlp2_step_workspace_words(B, D, H, C, ks)
        self.workspace = torch.zeros(words, dtype=torch.float32, device=self.device)
        self.max_steps = max_steps
        # loss | acc | step share one buffer so a caller can bring all three to the host with one copy
        self.out = torch.zeros(3 * max_steps, dtype=torch.float32, device=self.device)
        self.loss = self.out[:max_steps]
        self.acc = self.out[max_steps:2 * max_steps]
        self.step = self.out[2 * max_steps:].view(torch.int32)
        self._out_host = torch.zeros(3 * max_steps, dtype=torch.float32).pin_memory()
        self.slices = lib().dtfe_mlp2_step_grid(D, ks)
        with torch.cuda.device(self.device):
            self.grid = lib().dtfe_mlp2_step_ctas(B, D, H, ks)     # CTAs per launch (rows of the dbg stamp array)
        if self.grid < 0:
            check(1, "dtfe_mlp2_step_ctas")

    def results_host(self, n: int):
        """(loss[n], acc[n], step[n]) of the last launch as Python lists: one D2H copy, one synchronisation."""
        self._out_host.copy_(self.out, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        m = self.max_steps
        h = self._out_host
        return h[:n].tolist(), h[m:m + n].tolist(), h[2 * m:2 * m + n].view(torch.int32).tolist()

    def _launch_struct(self, W1, b1, W2, b2, global_step, lr, atomic, train, remote, out_offset=0):
        L = _Mlp2Launch()
        L.B, L.D, L.H, L.C, L.ks = self.B, self.D, self.H, self.C, self.ks
        L.W1, L.b1, L.W2, L.b2 = W1.data_ptr(), b1.data_ptr(), W2.data_ptr(), b2.data_ptr()
        L.global_step = global_step.data_ptr()
        L.workspace = self.workspace.data_ptr()
        L.loss_out = self.loss.data_ptr() + 4 * out_offset
        L.acc_out = self.acc.data_ptr() + 4 * out_offset
        L.step_out = self.step.data_ptr() + 4 * out_offset
        L.lr, L.atomic, L.train, L.remote = lr, int(atomic), int(train), int(remote)
        return L

    def train_host(self, W1, b1, W2, b2, global_step, x_host, y_host, loss_host, lr: float, atomic: bool = True,
                   remote: bool = True, steps_per_launch: int = 1):
        """Native host-fed loop: ``x_host``/``y_host`` are lists of pinned CPU tensors (one batch per step),
        ``loss_host`` a pinned float32 tensor with one slot per step."""
        if getattr(self, "_feed", None) is None or self._feed_spl != steps_per_launch:
            if getattr(self, "_feed", None) is not None:
                torch.cuda.synchronize(self.device)
                lib().dtfe_hostfeed_destroy(self._feed)
            h = _P()
            with torch.cuda.device(self.device):
                check(lib().dtfe_hostfeed_create(self.B, self.D, steps_per_launch, ctypes.byref(h)), "dtfe_hostfeed_create")
            self._feed, self._feed_spl = h, steps_per_launch
        n = len(x_host)
        if len(y_host) != n:
            raise ValueError("x_host and y_host must have one entry per step")
        # Validate and resolve every DISTINCT tensor once (is_pinned() is a driver query of about a microsecond;
        # a feed loop cycles through a ring of staging tensors, so per-entry checks would cost more than the
        # driver calls of the loop itself).
        xptr, yptr = {}, {}
        for x in x_host:
            if id(x) not in xptr:
                if not (x.is_pinned() and x.dtype == torch.float32 and x.is_contiguous()
                        and tuple(x.shape) == (self.B, self.D)):
                    raise TypeError("host batches must be pinned, contiguous float32 [B,D] tensors")
                xptr[id(x)] = x.data_ptr()
        for y in y_host:
            if id(y) not in yptr:
                if not (y.is_pinned() and y.dtype == torch.int32 and y.is_contiguous() and y.numel() == self.B):
                    raise TypeError("host labels must be pinned, contiguous int32 [B] tensors")
                yptr[id(y)] = y.data_ptr()
        if not loss_host.is_pinned() or loss_host.numel() < n or loss_host.dtype != torch.float32:
            raise TypeError("loss_host must be a pinned float32 tensor with one element per step")
        XP = (_P * n)(*[xptr[id(x)] for x in x_host])
        YP = (_P * n)(*[yptr[id(y)] for y in y_host])
        L = self._launch_struct(W1, b1, W2, b2, global_step, lr, atomic, True, remote)
        check(lib().dtfe_mlp2_train_host(self._feed, ctypes.byref(L), XP, YP, n, loss_host.data_ptr(), _stream()),
              "dtfe_mlp2_train_host")

    def __del__(self):
        try:
            if getattr(self, "_feed", None) is not None:
                lib().dtfe_hostfeed_destroy(self._feed)
        except Exception:
            pass

    def run(self, W1, b1, W2, b2, global_step, X, labels, lr: float, nsteps: int = 1, batch_idx=None,
            rows_per_step: int = 0, atomic: bool = True, train: bool = True, out_offset: int = 0,
            remote: bool = True, dbg=None, grads=None, cursor=None, sync_exchange=None, outs=None, tc: bool = False):
        """``tc``: run the tensor-core form of the step (``dtfe_mlp2_step_tc``: tcgen05 kind::tf32 on one 8-CTA
        cluster) instead of the fp32 SIMT kernel; asynchronous lossless training only.
        ``outs``: (loss, acc, step) device tensors (float32, float32, int32; >= nsteps elements each) that receive
        the per-step results instead of this object's own ring."""
        assert outs is not None or nsteps + out_offset <= self.max_steps
        L = _Mlp2Launch()
        L.B, L.D, L.H, L.C, L.ks = self.B, self.D, self.H, self.C, self.ks
        L.W1, L.b1, L.W2, L.b2 = W1.data_ptr(), b1.data_ptr(), W2.data_ptr(), b2.data_ptr()
        L.global_step = global_step.data_ptr()
        L.X, L.ldx = X.data_ptr(), X.stride(0)
        L.labels = labels.data_ptr()
        L.batch_idx = _ptr(batch_idx)
        L.rows_per_step = rows_per_step
        L.workspace = self.workspace.data_ptr()
        L.loss_out = self.loss.data_ptr() + 4 * out_offset
        L.acc_out = self.acc.data_ptr() + 4 * out_offset
        L.step_out = self.step.data_ptr() + 4 * out_offset
        if outs is not None:
            lo, ao, so = outs
            assert lo.numel() >= nsteps and ao.numel() >= nsteps and so.numel() >= nsteps and so.dtype == torch.int32
            L.loss_out, L.acc_out, L.step_out = lo.data_ptr(), ao.data_ptr(), so.data_ptr()
        L.lr, L.nsteps, L.atomic, L.train = lr, nsteps, int(atomic), int(train)
        L.remote = int(remote)
        L.dbg = _ptr(dbg)
        if grads is not None:
            L.gW1, L.gb1, L.gW2, L.gb2 = (g.data_ptr() for g in grads)
        L.cursor = _ptr(cursor)
        if sync_exchange is not None:
            L.sync_exchange = ctypes.addressof(sync_exchange)
        if tc:
            check(lib().dtfe_mlp2_step_tc(ctypes.byref(L), _stream()), "dtfe_mlp2_step_tc")
        else:
            check(lib().dtfe_mlp2_step(ctypes.byref(L), _stream()), "dtfe_mlp2_step")
