"""ncu driver: one 20-step launch of the fused step kernel in its lean form (asynchronous training) and one in its
full form with the synchronous exchange as phase D (pair transport, world = 1: every protocol step runs, the peer
stores and polls go to this GPU's own receive buffer)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from distributedtensorflowexample_b200.models.mlp import reference_mlp_spec
from distributedtensorflowexample_b200.parallel.spmd import build_spmd_store
from distributedtensorflowexample_b200.parallel.sync import SyncGroup
from distributedtensorflowexample_b200.train.gpu_step import DeviceDataset, FusedMlp2Stepper
from distributedtensorflowexample_b200.data import synthetic_arrays
dev = torch.device("cuda:0"); torch.cuda.set_device(0)
spec = reference_mlp_spec()
model, store = build_spmd_store(spec, 0, 1, dev)
names = [v.name for v in model.global_trainable]
x, y = synthetic_arrays(55000, seed=0)
ds = DeviceDataset(x, y.astype(np.int32), dev, seed=0)
n = 20
a = FusedMlp2Stepper(spec, 100, [store.tensor(k) for k in names], store.tensor("global/global_step"), 0.001, dev,
                     max_steps_per_launch=n, remote=False)
sg = SyncGroup(store, names, 0, 1, timeout_s=5.0, mode="allreduce_ll"); sg.refresh_replicas()
s = FusedMlp2Stepper(spec, 100, sg.home_params(), sg.home("global/global_step"), 0.001, dev, max_steps_per_launch=n,
                     remote=False, grad_buffers=sg.grad_views({k: tuple(store.tensor(k).shape) for k in names}))
s.sync_exchange = sg.exchange_struct(0.001)
for rep in range(3):
    for st in (a, s):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        idx = ds.next_indices(100, n)
        e0.record(); st.run(ds.images, ds.labels, nsteps=n, batch_idx=idx); e1.record(); torch.cuda.synchronize()
        print(f"{'sync (full, pair transport, W=1)' if st is s else 'async (lean)':34s} {e0.elapsed_time(e1) * 1e3 / n:6.2f} us/step")
