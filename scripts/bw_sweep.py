"""ps push/pull bandwidth sweep, 4 KB - 1 GB shards, vs NCCL (BASELINE.json config 5).

Run under torchrun with N = 2/4/8 ranks (one per GPU):
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 scripts/bw_sweep.py [--max-mb 1024]

Patterns (every rank is a worker; rank r's "ps shard" lives in rank r's HBM):
  push_ring : rank r pushes (red.add.sys, fused scale = the SGD apply) into the shard of rank r+1   -> per-link rate
  pull_ring : rank r pulls (L1-bypassing vector loads) from the shard of rank r+1                     -> per-link rate
  push_ps   : every rank pushes into rank 0's shard (the reference's star topology, many-to-one)
  pull_ps   : every rank pulls rank 0's shard (one-to-many)
NCCL baselines for the same bytes: send/recv ring (p2p), reduce-to-0 (push_ps), broadcast-from-0 (pull_ps).
Times are CUDA events on the launching stream, max over ranks; GB/s counts payload bytes per rank.
"""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from distributedtensorflowexample_b200.ops import cuda_ops as ops


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-mb", type=int, default=1024)
    ap.add_argument("--iters", type=int, default=10)
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    max_bytes = args.max_mb << 20
    # one IPC-exportable shard per rank, opened by everyone
    ptr = ops.shard_alloc(max_bytes)
    handles = [None] * world
    dist.all_gather_object(handles, ops.ipc_export(ptr))
    ptrs = [ptr if r == rank else ops.ipc_open(h) for r, h in enumerate(handles)]
    shards = [ops.tensor_from_ptr(p, (max_bytes // 4,), torch.float32, dev) for p in ptrs]
    grad = torch.ones(max_bytes // 4, device=dev) * 1e-6
    local_buf = torch.empty(max_bytes // 4, device=dev)
    nxt = (rank + 1) % world

    def timed(fn, iters):
        for _ in range(3):
            fn()
        dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    size = 4096
    rows = []
    while size <= max_bytes:
        n = size // 4
        it = args.iters if size >= (1 << 20) else 50
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
        if rank == 0:
            print(json.dumps(r), flush=True)
        size *= 4
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
