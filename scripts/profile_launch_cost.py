"""Where the fixed cost of one launch of the fused MNIST step kernel goes.

For nsteps in a sweep, three clocks around ONE launch of `nsteps` steps:
  idle     start event recorded on an idle stream (what bench.py r1 did): includes the host's own time between the
           event record and the kernel reaching the GPU;
  queued   the same, but queued behind a 256 MB fill kernel, so the launch is already waiting when the event fires;
  kernel   %globaltimer inside the kernel: first stamp of the first step (min over CTAs) .. last stamp of the last
           step (max over CTAs), plus the period of the first steps vs the steady state.
"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from distributedtensorflowexample_b200.models.mlp import reference_mlp_spec
from distributedtensorflowexample_b200.parallel.spmd import build_spmd_store
from distributedtensorflowexample_b200.train.gpu_step import DeviceDataset
from distributedtensorflowexample_b200.train.trainer import GpuTrainer
from distributedtensorflowexample_b200.data import synthetic_arrays
from distributedtensorflowexample_b200.ops import cuda_ops as ops

dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
spec = reference_mlp_spec()
model, store = build_spmd_store(spec, 0, 1, dev)
names = [v.name for v in model.global_trainable]
x, y = synthetic_arrays(55000, seed=0)
ds = DeviceDataset(x, y.astype(np.int32), dev, seed=0)
tr = GpuTrainer(spec, 100, 0.001, [store.tensor(n) for n in names], store.tensor("global/global_step"), dev,
                steps_per_launch=512, remote=False)
flush = torch.empty(64 << 20, dtype=torch.int32, device=dev)
G = tr.fused.k.grid
med = lambda v: sorted(v)[len(v) // 2]
tr.run_indices(ds, ds.next_indices(100, 50), 50)
torch.cuda.synchronize()
print(f"{'nsteps':>6} {'idle us':>9} {'queued us':>10} {'kernel us':>10} {'idle/step':>10} {'queued/step':>11} {'kernel/step':>11}")
rows = []
for n in (1, 2, 5, 10, 20, 50, 100, 200, 500):
    idle, queued, kern, first = [], [], [], []
    for rep in range(9):
        idx = ds.next_indices(100, n)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record(); tr.run_indices(ds, idx, n); e1.record(); torch.cuda.synchronize()
        idle.append(e0.elapsed_time(e1) * 1e3)
        idx = ds.next_indices(100, n)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ops.l2_flush(flush)
        e0.record(); tr.run_indices(ds, idx, n); e1.record(); torch.cuda.synchronize()
        queued.append(e0.elapsed_time(e1) * 1e3)
        dbg = torch.zeros(n * G * 16, dtype=torch.int64, device=dev)
        tr.fused.dbg = dbg
        idx = ds.next_indices(100, n)
        ops.l2_flush(flush)
        tr.run_indices(ds, idx, n)
        torch.cuda.synchronize()
        tr.fused.dbg = None
        t = dbg.view(n, G, 16).double().cpu().numpy()
        last = t[n - 1][:, [5, 7]].max()
        kern.append((last - t[0, :, 0].min()) / 1e3)
        if n >= 5:
            first.append([(t[i + 1, :, 0].min() - t[i, :, 0].min()) / 1e3 for i in range(4)])
    rows.append((n, med(idle), med(queued), med(kern)))
    print(f"{n:6d} {med(idle):9.1f} {med(queued):10.1f} {med(kern):10.1f} {med(idle)/n:10.2f} {med(queued)/n:11.2f} {med(kern)/n:11.2f}")
    if first:
        f = np.median(np.array(first), axis=0)
        print("        period of steps 0..3 inside the kernel:", " ".join(f"{v:.2f}" for v in f), "us")
# fixed cost = intercept of a line through the two largest points
(n1, i1, q1, k1), (n2, i2, q2, k2) = rows[-2], rows[-1]
for name, a, b in (("idle", i1, i2), ("queued", q1, q2), ("kernel", k1, k2)):
    slope = (b - a) / (n2 - n1)
    print(f"{name:7s}: {slope:6.2f} us/step steady, fixed {b - slope * n2:6.1f} us per launch")
store.close()
