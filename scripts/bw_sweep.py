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
from distributedtensorflowexample_b200.bench_bw import sweep


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
    sweep(rank, world, dev, args.max_mb, args.iters, emit=lambda r: print(json.dumps(r), flush=True))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
