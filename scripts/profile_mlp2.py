"""Tiny driver for ncu: a few launches of the fused MNIST step kernel (see profiles/README.md)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from distributedtensorflowexample_b200.models.mlp import reference_mlp_spec
from distributedtensorflowexample_b200.parallel.spmd import build_spmd_store
from distributedtensorflowexample_b200.train.gpu_step import DeviceDataset
from distributedtensorflowexample_b200.train.trainer import GpuTrainer
from distributedtensorflowexample_b200.data import synthetic_arrays

nsteps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
launches = int(sys.argv[2]) if len(sys.argv) > 2 else 4
rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=dev)
spec = reference_mlp_spec()
model, store = build_spmd_store(spec, rank, world, dev)
names = [v.name for v in model.global_trainable]
x, y = synthetic_arrays(55000, seed=0)
ds = DeviceDataset(x, y.astype(np.int32), dev, seed=0)
tr = GpuTrainer(spec, 100, 0.001, [store.tensor(n) for n in names], store.tensor("global/global_step"), dev,
                steps_per_launch=nsteps, remote=('--remote' in sys.argv) or world > 1)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for i in range(launches):
    idx = ds.next_indices(100, nsteps)
    e0.record()
    tr.run_indices(ds, idx, nsteps)
    e1.record()
    torch.cuda.synchronize()
    if rank == world - 1:
        print(f"[rank {rank}] launch {i}: {e0.elapsed_time(e1) * 1e3 / nsteps:.2f} us/step")

if world > 1:
    dist.barrier()
if "--stamps" in sys.argv and rank == world - 1:
    G = tr.fused.k.grid
    dbg = torch.zeros(nsteps * G * 16, dtype=torch.int64, device=dev)
    tr.fused.dbg = dbg
    idx = ds.next_indices(100, nsteps)
    tr.run_indices(ds, idx, nsteps)
    torch.cuda.synchronize()
    t = dbg.view(nsteps, G, 16).double().cpu().numpy()[2:]      # skip the first steps
    def seg(name, a, b, rows=slice(None)):
        d = (t[:, rows, b] - t[:, rows, a]) / 1e3
        print(f"{name:34s} mean {d.mean():6.2f} us  max-over-CTAs {d.max(axis=1).mean():6.2f}  min {d.min(axis=1).mean():6.2f}")
    act = slice(0, min(100, G))                                 # CTAs that own batch rows in phase B
    sl = slice(0, tr.fused.k.slices)                            # CTAs that own a W1 slice
    seg("A: load+compute+store", 0, 1, sl)
    seg("barrier1", 1, 2)
    seg("B total", 2, 3)
    seg("  B (row CTAs) loads+reduce", 2, 8, act)
    seg("  B (row CTAs) stage+sigmoid", 8, 9, act)
    seg("  B (row CTAs) logits", 9, 10, act)
    seg("  B (row CTAs) softmax/xent", 10, 11, act)
    seg("  B (row CTAs) dA+store", 11, 3, act)
    seg("barrier2", 3, 4)
    seg("C total (slice CTAs)", 4, 5, sl)
    seg("  C wait for dA chunk 0", 4, 12, sl)
    seg("  C dW1 (dA streamed in)", 12, 13, sl)
    seg("  C wait W1 fetch + push", 13, 5, sl)
    hp = slice(tr.fused.k.slices, G)
    if G > tr.fused.k.slices:
        seg("C helpers: wait dZ", 4, 12, hp)
        seg("C helpers: small grads", 12, 14, hp)
        seg("C helpers: tail", 14, 5, hp)
    seg("  C closing __syncthreads", 5, 7)
    bt = ((t[:, :, 3] - t[:, :, 2]) / 1e3).mean(axis=0)          # phase B duration per CTA
    worst = bt.argsort()[::-1][:6]
    print("slowest CTAs in phase B:", ", ".join(f"cta {int(c)}: {bt[c]:.2f} us" for c in worst))
    ct = ((t[:, :, 5] - t[:, :, 4]) / 1e3).mean(axis=0)
    worst = ct.argsort()[::-1][:6]
    print("slowest CTAs in phase C:", ", ".join(f"cta {int(c)}: {ct[c]:.2f} us" for c in worst))
    step = (t[1:, :, 0] - t[:-1, :, 0]) / 1e3
    print(f"step period             {step.mean():6.2f} us")

if world > 1:
    dist.barrier()
    dist.destroy_process_group()
