"""ps push/pull bandwidth sweep behind ``bench.py --config bw`` (BASELINE.json config 5): 4 KB - 1 GB shards at
N = 2/4/8 GPUs, this framework's push (red.add with the SGD scale fused = the apply) and pull (L1-bypassing vector
loads) over peer memory against the NCCL collectives that move the same bytes.

Patterns (every rank is a worker; rank r's "ps shard" lives in rank r's HBM):
  push_ring : rank r pushes into the shard of rank r+1            -> per-link rate
  pull_ring : rank r pulls from the shard of rank r+1             -> per-link rate
  push_ps   : every rank pushes into rank 0's shard (the reference's star topology, many-to-one)
  pull_ps   : every rank pulls rank 0's shard (one-to-many)
NCCL baselines for the same bytes: send/recv ring (p2p), reduce-to-0 (push_ps), broadcast-from-0 (pull_ps).
Times are CUDA events on the launching stream, max over ranks; GB/s counts payload bytes per rank.
"""
from __future__ import annotations

import json

import torch
import torch.distributed as dist

from .ops import cuda_ops as ops


def sweep(rank: int, world: int, dev, max_mb: int = 1024, iters: int = 10, emit=None):
    max_bytes = max_mb << 20
    # one IPC-exportable shard per rank, opened by everyone
    ptr = ops.shard_alloc(max_bytes)
    handles = [None] * world
    dist.all_gather_object(handles, ops.ipc_export(ptr))
    ptrs = [ptr if r == rank else ops.ipc_open(h) for r, h in enumerate(handles)]
    shards = [ops.tensor_from_ptr(p, (max_bytes // 4,), torch.float32, dev) for p in ptrs]
    grad = torch.ones(max_bytes // 4, device=dev) * 1e-6
    local_buf = torch.empty(max_bytes // 4, device=dev)
    nxt = (rank + 1) % world

    def timed(fn, it):
        for _ in range(3):
            fn()
        dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(it):
            fn()
        e1.record(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / it], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    size = 4096
    rows = []
    while size <= max_bytes:
        n = size // 4
        it = iters if size >= (1 << 20) else 50
        g, lb = grad[:n], local_buf[:n]
        r = {"bytes": size, "ranks": world}
        r["push_ring_us"] = timed(lambda: ops.push_sgd(shards[nxt][:n], g, -0.001, atomic=True), it) * 1e3
        r["pull_ring_us"] = timed(lambda: ops.pull(lb, shards[nxt][:n]), it) * 1e3
        r["push_ps_us"] = timed(lambda: ops.push_sgd(shards[0][:n], g, -0.001, atomic=True), it) * 1e3
        r["pull_ps_us"] = timed(lambda: ops.pull(lb, shards[0][:n]), it) * 1e3

        def nccl_ring():
            ops_ = [dist.P2POp(dist.isend, g, nxt), dist.P2POp(dist.irecv, lb, (rank - 1) % world)]
            for w in dist.batch_isend_irecv(ops_):
                w.wait()
        r["nccl_sendrecv_ring_us"] = timed(nccl_ring, it) * 1e3
        r["nccl_reduce_us"] = timed(lambda: dist.reduce(lb, dst=0), it) * 1e3
        r["nccl_broadcast_us"] = timed(lambda: dist.broadcast(lb, src=0), it) * 1e3
        for k in list(r):
            if k.endswith("_us"):
                r[k.replace("_us", "_GBs")] = round(size / (r[k] * 1e-6) / 1e9, 2)
                r[k] = round(r[k], 2)
        rows.append(r)
        if emit is not None and rank == 0:
            emit(r)
        size *= 4
    dist.barrier()
    for r_, p in enumerate(ptrs):
        if r_ != rank:
            ops.ipc_close(p)
    dist.barrier()
    ops.shard_free(ptr)
    return rows


def run(args, rank: int, world: int, dev) -> int:
    if world < 2:
        if rank == 0:
            print(json.dumps({"metric": "ps push/pull bandwidth sweep", "error": "needs >= 2 GPUs (torchrun, one rank per GPU)"}))
        return 0
    rows = sweep(rank, world, dev, max_mb=getattr(args, "bw_max_mb", 1024))
    if rank == 0:
        big = rows[-1]
        print(json.dumps({"metric": "ps push/pull bandwidth sweep 4 KB - %d MB (GB/s per rank, device-timed, max over ranks)" % (rows[-1]["bytes"] >> 20),
                          "value": big["push_ring_GBs"], "unit": "GB/s (push_ring at the largest size)", "n_gpus": world,
                          "higher_is_better": True, "impl": "ours", "vs_baseline": round(big["push_ring_GBs"] / max(big["nccl_sendrecv_ring_GBs"], 1e-9), 3),
                          "vs_baseline_denominator": "NCCL send/recv ring, same bytes", "rows": rows}), flush=True)
    dist.barrier()
    dist.destroy_process_group()
    return 0
