"""Time-to-target test accuracy of the NCCL baseline (baseline/nccl_ps.py: flat all-reduce, whole step in a CUDA
graph) on the same data, model, batch, learning rate and target as scripts/tta_launcher.py.  Run under torchrun.

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 scripts/tta_nccl.py
"""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from baseline.nccl_ps import NcclPsWorker
from distributedtensorflowexample_b200.data import find_data_dir, read_data_sets

ap = argparse.ArgumentParser()
ap.add_argument("--target", type=float, default=0.88)
ap.add_argument("--lr", type=float, default=0.1)
ap.add_argument("--eval-every", type=int, default=500)
ap.add_argument("--max-steps", type=int, default=60000)
ap.add_argument("--repeats", type=int, default=3)
args = ap.parse_args()
rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
B = 100
for rep in range(args.repeats):
    mn = read_data_sets(find_data_dir(), one_hot=False, seed=rep)
    Xtr = torch.tensor(mn.train.images, device=dev); Ytr = torch.tensor(mn.train.label_ids, device=dev).int()
    Xte = torch.tensor(mn.test.images, device=dev); Yte = torch.tensor(mn.test.label_ids, device=dev)
    n = Xtr.shape[0]
    wk = NcclPsWorker((784, 100, 10), B, args.lr, rank, world, dev, torch.float32, flat=True)
    gen = torch.Generator(device=dev).manual_seed(1000 * rep + rank)
    xs, ys = torch.zeros(B, 784, device=dev), torch.zeros(B, device=dev, dtype=torch.int32)
    perm = torch.randperm(n, device=dev, generator=gen)
    cursor = torch.zeros(1, device=dev, dtype=torch.long)
    ar = torch.arange(B, device=dev)
    nb = n // B

    def step():
        rows = perm[(cursor % nb) * B + ar]
        torch.index_select(Xtr, 0, rows, out=xs); torch.index_select(Ytr, 0, rows, out=ys)
        cursor.add_(1)
        wk.step(xs, ys)

    s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(3):
            step()
    torch.cuda.current_stream().wait_stream(s); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        step()
    # restart from a fresh N(0,1) initialisation (same on every rank), like the product arm
    gi = torch.Generator(device=dev).manual_seed(rep)
    wk.flat_p.zero_()
    wk.params[0].copy_(torch.randn(784, 100, device=dev, generator=gi)); wk.params[2].copy_(torch.randn(100, 10, device=dev, generator=gi))
    cursor.zero_()

    def evaluate():
        h = torch.sigmoid(Xte @ wk.params[0] + wk.params[1])
        return float(((h @ wk.params[2] + wk.params[3]).argmax(1) == Yte).float().mean())

    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter(); steps = 0; acc = 0.0
    while steps < args.max_steps:
        for _ in range(args.eval_every):
            g.replay()
        steps += args.eval_every
        acc = evaluate()
        if acc >= args.target:
            break
    torch.cuda.synchronize()
    if rank == 0:
        print(json.dumps({"impl": "nccl-ps graph (sync all-reduce)", "workers": world, "seed": rep, "target": args.target,
                          "lr": args.lr, "eval_every": args.eval_every, "reached": bool(acc >= args.target),
                          "seconds": round(time.perf_counter() - t0, 4), "global_step": steps, "test_accuracy": round(acc, 4)}), flush=True)
if world > 1:
    dist.barrier(); dist.destroy_process_group()
