"""Per-phase %globaltimer stamps of the tensor-core fused step (mlp2_step_tc.cu) next to the SIMT kernel's period."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from distributedtensorflowexample_b200.models.mlp import reference_mlp_spec
from distributedtensorflowexample_b200.parallel.spmd import build_spmd_store
from distributedtensorflowexample_b200.train.gpu_step import DeviceDataset, FusedMlp2Stepper
from distributedtensorflowexample_b200.data import synthetic_arrays
from distributedtensorflowexample_b200.ops import cuda_ops as ops

nsteps = int(sys.argv[1]) if len(sys.argv) > 1 else 200
dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
spec = reference_mlp_spec()
model, store = build_spmd_store(spec, 0, 1, dev)
names = [v.name for v in model.global_trainable]
x, y = synthetic_arrays(55000, seed=0)
ds = DeviceDataset(x, y.astype(np.int32), dev, seed=0)
args = ([store.tensor(n) for n in names], store.tensor("global/global_step"), 0.001, dev)
flush = torch.empty(64 << 20, dtype=torch.int32, device=dev)
for tc in (False, True):
    st = FusedMlp2Stepper(spec, 100, *args, max_steps_per_launch=nsteps, remote=False, tc=tc)
    st.run(ds.images, ds.labels, nsteps=nsteps, batch_idx=ds.next_indices(100, nsteps))
    torch.cuda.synchronize()
    ts = []
    for rep in range(9):
        idx = ds.next_indices(100, nsteps)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ops.l2_flush(flush)
        e0.record(); st.run(ds.images, ds.labels, nsteps=nsteps, batch_idx=idx); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3 / nsteps)
    print(f"{'tcgen05 tf32 (8-CTA cluster)' if tc else 'fp32 SIMT (148-CTA cooperative grid)'}: "
          f"{sorted(ts)[len(ts)//2]:.2f} us/step (median of 9 launches of {nsteps} steps), loss {st.loss[nsteps-1].item():.4f}")
    if tc:
        dbg = torch.zeros(nsteps * 8 * 16, dtype=torch.int64, device=dev)
        st.dbg = dbg
        st.run(ds.images, ds.labels, nsteps=nsteps, batch_idx=ds.next_indices(100, nsteps))
        torch.cuda.synchronize()
        t = dbg.view(nsteps, 8, 16).double().cpu().numpy()[2:]
        def seg(name, a, b):
            d = (t[:, :, b] - t[:, :, a]) / 1e3
            print(f"  {name:46s} mean {d.mean():6.2f} us   max-over-CTAs {d.max(axis=1).mean():6.2f}   min {d.min(axis=1).mean():6.2f}")
        seg("wait X/W1 + 13 forward MMAs", 0, 1)
        seg("issue W1 pull (TMA) + tcgen05.ld", 1, 9)
        seg("partial -> scratch", 9, 2)
        seg("cluster barrier 1", 2, 3)
        seg("B: issue loads, small params, __syncthreads", 3, 10)
        seg("B: sum 8 partials + bias + sigmoid", 10, 11)
        seg("B: logits + softmax-xent", 11, 12)
        seg("B: dA", 12, 4)
        seg("cluster barrier 2", 4, 5)
        seg("C: stage dA/dZ (cp.async) + issue next X", 5, 13)
        seg("C: issue 13 backward MMAs", 13, 6)
        seg("C: small gradients (under the MMAs)", 6, 14)
        seg("C: wait for the MMAs", 14, 7)
        seg("C: tcgen05.ld -> red.add.v4 push + own delta", 7, 8)
        step = (t[1:, :, 0] - t[:-1, :, 0]) / 1e3
        print(f"  step period {step.mean():.2f} us")
store.close()
