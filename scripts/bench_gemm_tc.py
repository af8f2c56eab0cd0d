"""Time the tcgen05 GEMM on the wide-MLP shapes against torch.matmul (cuBLAS) on the same operands."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributedtensorflowexample_b200.ops import gemm_tc as G
from distributedtensorflowexample_b200.ops import cuda_ops

def timeit(fn, iters=20, flush=None):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush is not None: cuda_ops.l2_flush(flush)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]

B, D0, H = 4096, 784, 8192
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
X = torch.rand(B, D0, device="cuda").bfloat16()
W1 = (torch.randn(D0, H, device="cuda") * 0.05).bfloat16()
W2 = (torch.randn(H, H, device="cuda") * 0.02).bfloat16()
b = torch.zeros(H, device="cuda").bfloat16()
H1 = torch.rand(B, H, device="cuda").bfloat16()
dA = (torch.randn(B, H, device="cuda") * 0.1).bfloat16()
out = torch.empty(B, H, device="cuda", dtype=torch.bfloat16)
rows = []
cases = [
    ("fwd1 X.W1 +bias+sigmoid", 2.0 * B * D0 * H, lambda tn: G.dense_fwd_tc(X, W1, b, 1, out, tile_n=tn), lambda: torch.sigmoid(X @ W1 + b)),
    ("fwd2 H1.W2 +bias+sigmoid", 2.0 * B * H * H, lambda tn: G.dense_fwd_tc(H1, W2, b, 1, out, tile_n=tn), lambda: torch.sigmoid(H1 @ W2 + b)),
    ("dX  dA.W2^T *sigmoid'", 2.0 * B * H * H, lambda tn: G.dense_bwd_dx_tc(dA, W2, H1, out, tile_n=tn), lambda: (dA @ W2.T) * H1 * (1 - H1)),
    ("dW2 H1^T.dA push(bf16)", 2.0 * B * H * H, lambda tn: G.dense_bwd_dw_push_tc(H1, dA, W2, -1e-6, tile_n=tn), lambda: W2.add_(H1.T @ dA, alpha=-1e-6)),
    ("dW1 X^T.dA push(bf16)", 2.0 * B * D0 * H, lambda tn: G.dense_bwd_dw_push_tc(X, dA, W1, -1e-6, tile_n=tn), lambda: W1.add_(X.T @ dA, alpha=-1e-6)),
]
for name, flops, ours, ref in cases:
    t_ref = timeit(ref, flush=flush)
    r = {"case": name, "torch_ms": round(t_ref, 4), "torch_tflops": round(flops / t_ref / 1e9, 1)}
    for tn in (128, 256, 512):
        t = timeit(lambda: ours(tn), flush=flush)
        r[f"ours_bn{tn}_ms"] = round(t, 4)
        r[f"ours_bn{tn}_tflops"] = round(flops / t / 1e9, 1)
    print(json.dumps(r), flush=True)
# plain matmul only (no epilogue) for the cuBLAS GEMM-only reference
t = timeit(lambda: torch.matmul(H1, W2), flush=flush)
print(json.dumps({"case": "torch.matmul H1.W2 only", "ms": round(t, 4), "tflops": round(2.0 * B * H * H / t / 1e9, 1)}))
