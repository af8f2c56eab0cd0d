"""Thin output layer (8192 -> 10, batch 4096, bf16): streaming kernels vs the 128-wide tensor-core tiles."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributedtensorflowexample_b200.ops import cuda_ops as ops, gemm_tc as G
M, K, C = 4096, 8192, 10
g = torch.Generator(device="cuda").manual_seed(0)
act = torch.rand(M, K, device="cuda", generator=g).bfloat16()
Wp = torch.zeros(K, 16, device="cuda", dtype=torch.bfloat16); Wp[:, :C] = (torch.randn(K, C, device="cuda", generator=g) * 0.02).bfloat16()
W = Wp[:, :C]
bias = torch.zeros(16, device="cuda", dtype=torch.bfloat16)[:C]
dZp = torch.zeros(M, 16, device="cuda", dtype=torch.bfloat16); dZp[:, :C] = (torch.randn(M, C, device="cuda", generator=g) / M).bfloat16()
dZ = dZp[:, :C]
out = torch.empty(M, 16, device="cuda")[:, :C]
dX = torch.empty(M, K, device="cuda", dtype=torch.bfloat16)
scratch = ops.thin_dw_scratch(M, K, "cuda")
flush = torch.empty(64 << 20, dtype=torch.int32, device="cuda")
def t(fn, n=10):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        ops.l2_flush(flush)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return sorted(ts)[len(ts) // 2]
rows = [("fwd  logits = act.W + b", lambda: ops.thin_fwd(act, W, bias, out), lambda: G.dense_fwd_tc(act, W, bias, 0, out), 2 * M * K),
        ("dX   (dZ.W^T)*h(1-h)", lambda: ops.thin_dx(dZ, W, act, dX), lambda: G.dense_bwd_dx_tc(dZ, W, act, dX), 4 * M * K),
        ("dW   W += a*act^T.dZ", lambda: ops.thin_dw_push(act, dZ, W, -1e-3, scratch), lambda: G.dense_bwd_dw_push_tc(act, dZ, W, -1e-3), 2 * M * K)]
for name, thin, tc, nbytes in rows:
    a, b = t(thin), t(tc)
    print(f"{name:26s} streaming {a:7.1f} us ({nbytes / a / 1e6:6.2f} TB/s of activation traffic)   tensor-core tiles {b:7.1f} us")
