"""ncu driver for the kernels whose operand or target lives in a PEER GPU's HBM (needs 2 GPUs, ONE process):
the shard is allocated on cuda:1 and mapped into this process with peer access, the kernels run on cuda:0.

  k_gemm_tc2  forward X.W + bias + sigmoid with W read by TMA from the peer shard          (pull fused with the GEMM)
  k_gemm_tc2  dW = X^T.dY pushed into the peer shard by TMA reduce-adds                     (push fused with the apply)
  k_push_sgd  flat push (red.add.v4) into the peer shard; k_pull flat pull                  (bandwidth sweep kernels)
"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributedtensorflowexample_b200.ops import cuda_ops as ops, gemm_tc as G
assert torch.cuda.device_count() >= 2
torch.cuda.set_device(0)
ops.enable_peer(0, 1)
M, K, N = 4096, 4096, 4096
g0 = torch.Generator(device="cuda:0").manual_seed(0)
X = torch.rand(M, K, device="cuda:0", generator=g0).bfloat16()
dY = (torch.randn(M, N, device="cuda:0", generator=g0) / M).bfloat16()
W_peer = (torch.randn(K, N, device="cuda:1") * 0.02).bfloat16()         # the "ps shard" on the other GPU
b = torch.zeros(N, device="cuda:0", dtype=torch.bfloat16)
out = torch.empty(M, N, device="cuda:0", dtype=torch.bfloat16)
flat_peer = torch.zeros(64 << 20, device="cuda:1")                        # 256 MB
grad = torch.full((64 << 20,), 1e-6, device="cuda:0")
local = torch.empty(64 << 20, device="cuda:0")
torch.cuda.synchronize(0); torch.cuda.synchronize(1)
def timed(name, fn, nbytes=None, flops=None):
    for _ in range(2): fn()
    torch.cuda.synchronize(0)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); fn(); e1.record(); torch.cuda.synchronize(0)
    us = e0.elapsed_time(e1) * 1e3
    extra = (f" {nbytes / us / 1e3:7.1f} GB/s over NVLink" if nbytes else "") + (f" {flops / us / 1e6:7.1f} TFLOP/s" if flops else "")
    print(f"{name:58s} {us:8.1f} us{extra}")
timed("forward GEMM, W (32 MB) pulled by TMA from the peer shard", lambda: G.dense_fwd_tc(X, W_peer, b, 1, out, tile_n=512),
      nbytes=K * N * 2, flops=2.0 * M * N * K)
timed("dW GEMM, pushed into the peer shard by TMA reduce-adds", lambda: G.dense_bwd_dw_push_tc(X, dY, W_peer, -1e-3, tile_n=512, tma_push=True),
      nbytes=K * N * 2, flops=2.0 * M * N * K)
timed("dW GEMM, pushed into the peer shard by red.add.v4.bf16x2", lambda: G.dense_bwd_dw_push_tc(X, dY, W_peer, -1e-3, tile_n=512),
      nbytes=K * N * 2, flops=2.0 * M * N * K)
timed("flat push 256 MB (red.add.v4.f32) into the peer shard", lambda: ops.push_sgd(flat_peer, grad, -1e-3), nbytes=256 << 20)
timed("flat pull 256 MB from the peer shard", lambda: ops.pull(local, flat_peer), nbytes=256 << 20)
