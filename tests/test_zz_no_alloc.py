"""`dtfe_mlp2_step_noalloc`: the fused training step for callers that are not allowed to allocate
device memory (its scratch is an array inside the library image).  Called here through the plain
C ABI, exactly as such a caller would, and compared with `dtfe_mlp2_step` fed a torch-allocated
workspace.  Collected last on purpose: if it fails under `-x`, everything else has already run."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

B, D, H, C, KS, NSTEPS, LR = 100, 784, 100, 10, 8, 3, 0.05


def _abi():
    from distributedtensorflowexample_b200.ops import cuda_ops
    lib = cuda_ops.lib()
    lib.dtfe_mlp2_step_noalloc.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
    lib.dtfe_mlp2_step_noalloc.restype = ctypes.c_int
    lib.dtfe_mlp2_step_noalloc_max_words.argtypes = []
    lib.dtfe_mlp2_step_noalloc_max_words.restype = ctypes.c_size_t
    return lib, cuda_ops._Mlp2Launch


class _Replica:
    """One independent copy of parameters, outputs and (for the reference path) a workspace."""

    def __init__(self, params, words):
        dev = "cuda:0"
        self.p = [torch.tensor(a, device=dev) for a in params]
        self.gs = torch.zeros(1, dtype=torch.int32, device=dev)
        self.loss = torch.zeros(NSTEPS, device=dev)
        self.acc = torch.zeros(NSTEPS, device=dev)
        self.step = torch.zeros(NSTEPS, dtype=torch.int32, device=dev)
        self.ws = torch.zeros(words, device=dev) if words else None

    def launch_struct(self, Launch, X, Y, idx, b=B, h=H):
        L = Launch()
        L.B, L.D, L.H, L.C, L.ks = b, D, h, C, KS
        L.W1, L.b1, L.W2, L.b2 = (t.data_ptr() for t in self.p)
        L.global_step = self.gs.data_ptr()
        L.X, L.ldx, L.labels = X.data_ptr(), X.stride(0), Y.data_ptr()
        L.batch_idx, L.rows_per_step = idx.data_ptr(), 0
        L.workspace = self.ws.data_ptr() if self.ws is not None else None
        L.loss_out, L.acc_out, L.step_out = self.loss.data_ptr(), self.acc.data_ptr(), self.step.data_ptr()
        L.lr, L.nsteps, L.atomic, L.train, L.remote = LR, NSTEPS, 1, 1, 1
        L.dbg = None
        return L


def test_noalloc_entry_matches_workspace_entry_twice():
    lib, Launch = _abi()
    rng = np.random.default_rng(42)
    params = [rng.normal(size=(D, H)).astype(np.float32), np.zeros(H, np.float32),
              rng.normal(size=(H, C)).astype(np.float32), np.zeros(C, np.float32)]
    X = torch.tensor(rng.random((1000, D), dtype=np.float32), device="cuda:0")
    Y = torch.tensor(rng.integers(0, C, 1000).astype(np.int32), device="cuda:0")
    words = lib.dtfe_mlp2_step_workspace_words(B, D, H, C, KS)
    assert words <= lib.dtfe_mlp2_step_noalloc_max_words()
    ref, new = _Replica(params, words), _Replica(params, 0)
    stream = torch.cuda.current_stream().cuda_stream
    for call in range(2):                       # two calls in a row share the built-in array
        idx = torch.tensor(rng.integers(0, 1000, NSTEPS * B).astype(np.int32), device="cuda:0")
        Lr, Ln = ref.launch_struct(Launch, X, Y, idx), new.launch_struct(Launch, X, Y, idx)
        assert lib.dtfe_mlp2_step(ctypes.byref(Lr), stream) == 0, lib.dtfe_last_error()
        assert lib.dtfe_mlp2_step_noalloc(ctypes.byref(Ln), stream) == 0, lib.dtfe_last_error()
        torch.cuda.synchronize()
        np.testing.assert_allclose(new.loss.cpu().numpy(), ref.loss.cpu().numpy(), rtol=2e-4, atol=2e-4)
        np.testing.assert_allclose(new.acc.cpu().numpy(), ref.acc.cpu().numpy(), atol=1e-6)
        assert new.step.cpu().tolist() == ref.step.cpu().tolist() == list(range(call * NSTEPS, (call + 1) * NSTEPS))
        for a, b in zip(new.p, ref.p):
            np.testing.assert_allclose(a.cpu().numpy(), b.cpu().numpy(), rtol=1e-3, atol=2e-4)
        assert np.isfinite(new.loss.cpu().numpy()).all()
    assert new.gs.item() == ref.gs.item() == 2 * NSTEPS


def test_noalloc_entry_refuses_shapes_beyond_its_array():
    lib, Launch = _abi()
    b, h = 128, 128                             # scratch for this shape is ~6.6 MiB > 4 MiB
    assert lib.dtfe_mlp2_step_workspace_words(b, D, h, C, KS) > lib.dtfe_mlp2_step_noalloc_max_words()
    params = [np.zeros((D, h), np.float32), np.zeros(h, np.float32), np.zeros((h, C), np.float32),
              np.zeros(C, np.float32)]
    rep = _Replica(params, 0)
    X = torch.zeros(b, D, device="cuda:0"); Y = torch.zeros(b, dtype=torch.int32, device="cuda:0")
    idx = torch.zeros(NSTEPS * b, dtype=torch.int32, device="cuda:0")
    L = rep.launch_struct(Launch, X, Y, idx, b=b, h=h)
    rc = lib.dtfe_mlp2_step_noalloc(ctypes.byref(L), torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    msg = lib.dtfe_last_error().decode()
    assert "scratch" in msg and "4 MiB" in msg
    torch.cuda.synchronize()
    assert rep.gs.item() == 0                   # nothing ran
