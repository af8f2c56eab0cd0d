"""ncu driver: the first-layer forward GEMM of the wide config (4096x784 . 784x8192 + bias + sigmoid, bf16)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributedtensorflowexample_b200.ops import gemm_tc as G
g = torch.Generator(device="cuda").manual_seed(0)
X = torch.rand(4096, 784, device="cuda", generator=g).bfloat16()
W = (torch.randn(784, 8192, device="cuda", generator=g) * 0.03).bfloat16()
b = torch.zeros(8192, device="cuda", dtype=torch.bfloat16)
out = torch.empty(4096, 8192, device="cuda", dtype=torch.bfloat16)
tile = int(sys.argv[1]) if len(sys.argv) > 1 else 0
for _ in range(4):
    G.dense_fwd_tc(X, W, b, 1, out, tile_n=tile)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); G.dense_fwd_tc(X, W, b, 1, out, tile_n=tile); e1.record(); torch.cuda.synchronize()
print(f"fwd1 tile_n={tile}: {e0.elapsed_time(e1) * 1e3:.1f} us")
