"""Smallest driver for compute-sanitizer: a few steps of the fused step kernel (SIMT and tensor-core forms), the
all-reduce exchange with world = 1, one tcgen05 GEMM, the thin kernels and the push/pull/Adam kernels."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from distributedtensorflowexample_b200.models.mlp import reference_mlp_spec
from distributedtensorflowexample_b200.parallel.spmd import build_spmd_store
from distributedtensorflowexample_b200.parallel.sync import SyncGroup
from distributedtensorflowexample_b200.train.gpu_step import FusedMlp2Stepper
from distributedtensorflowexample_b200.data import synthetic_arrays
from distributedtensorflowexample_b200.ops import cuda_ops as ops, gemm_tc as G
dev = torch.device("cuda:0"); torch.cuda.set_device(0)
which = sys.argv[1] if len(sys.argv) > 1 else "all"
spec = reference_mlp_spec()
model, store = build_spmd_store(spec, 0, 1, dev)
names = [v.name for v in model.global_trainable]
x, y = synthetic_arrays(500, seed=0)
X = torch.tensor(x, device=dev); Y = torch.tensor(y.astype(np.int32), device=dev)
idx = torch.arange(300, device=dev, dtype=torch.int32)
args = ([store.tensor(n) for n in names], store.tensor("global/global_step"), 0.05, dev)
if which in ("all", "simt"):
    FusedMlp2Stepper(spec, 100, *args, max_steps_per_launch=3).run(X, Y, nsteps=3, batch_idx=idx)
if which in ("all", "tc"):
    FusedMlp2Stepper(spec, 100, *args, max_steps_per_launch=3, tc=True).run(X, Y, nsteps=3, batch_idx=idx)
if which in ("all", "sync"):
    for mode in ("allreduce", "allreduce_ll"):
        sg = SyncGroup(store, names, 0, 1, timeout_s=5.0, mode=mode); sg.refresh_replicas()
        shapes = {n: tuple(store.tensor(n).shape) for n in names}
        st = FusedMlp2Stepper(spec, 100, sg.home_params(), sg.home("global/global_step"), 0.05, dev, max_steps_per_launch=3,
                              remote=False, grad_buffers=sg.grad_views(shapes))
        st.sync_exchange = sg.exchange_struct(0.05)
        st.run(X, Y, nsteps=3, batch_idx=idx)
if which in ("all", "gemm"):
    G.smoke()
    a = torch.rand(64, 256, device=dev).bfloat16(); w = torch.zeros(256, 16, device=dev, dtype=torch.bfloat16)[:, :10]
    out = torch.empty(64, 16, device=dev)[:, :10]
    ops.thin_fwd(a, w, None, out)
    t = torch.zeros(4096, device=dev); ops.push_sgd(t, torch.ones(4096, device=dev), -1.0)
torch.cuda.synchronize()
print("sanitize_small ok", which)
