"""ncu driver: a few launches of the 2-CTA tcgen05 GEMM with the push epilogue (dW2 shape of the wide MLP)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributedtensorflowexample_b200.ops import gemm_tc as G
B, H = 4096, 8192
H1 = torch.rand(B, H, device="cuda").bfloat16()
dA = (torch.randn(B, H, device="cuda") * 0.1).bfloat16()
W2 = (torch.randn(H, H, device="cuda") * 0.02).bfloat16()
b = torch.zeros(H, device="cuda").bfloat16()
out = torch.empty(B, H, device="cuda", dtype=torch.bfloat16)
for _ in range(3):
    G.dense_bwd_dw_push_tc(H1, dA, W2, -1e-6)          # push epilogue (red.add into the shard)
    G.dense_fwd_tc(H1, W2, b, 1, out)                   # pull + bias + sigmoid
torch.cuda.synchronize()
print("ok")
