"""Run under torchrun with >= 2 GPUs (spawned by tests/test_gpu_multi.py):
  async : 2 workers push lossless red.add updates into shards on both GPUs - conservation check;
  sync  : reduce-scatter+apply over peer memory == single-process large-batch SGD."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

from distributedtensorflowexample_b200.data import synthetic_arrays
from distributedtensorflowexample_b200.models import oracle
from distributedtensorflowexample_b200.models.mlp import reference_mlp_spec
from distributedtensorflowexample_b200.parallel.spmd import build_spmd_store
from distributedtensorflowexample_b200.parallel.sync import SyncGroup
from distributedtensorflowexample_b200.train.gpu_step import LayerwiseStepper, FusedMlp2Stepper


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=dev)
    spec = reference_mlp_spec()
    B, lr = 100, 0.05
    x, y = synthetic_arrays(4000, seed=7)
    X = torch.tensor(x, device=dev); Y = torch.tensor(y.astype(np.int32), device=dev)

    # ---------------- sync: must equal one worker with the concatenated batch -----------------
    model, store = build_spmd_store(spec, rank, world, dev, seed=3)
    names = [v.name for v in model.global_trainable]
    p0 = [store.read(n).astype(np.float64) for n in names]
    sg = SyncGroup(store, names, rank, world)
    shapes = {n: tuple(store.tensor(n).shape) for n in names}
    st = LayerwiseStepper(spec, B, [store.tensor(n) for n in names], store.tensor("global/global_step"), lr, dev,
                          grad_buffers=sg.grad_views(shapes))
    nsteps = 5
    for s in range(nsteps):
        rows = slice((s * world + rank) * B, (s * world + rank + 1) * B)
        st.run(X[rows], Y[rows])
        sg.reduce_scatter_apply(lr)
    torch.cuda.synchronize(); dist.barrier()
    ref = [p.copy() for p in p0]
    for s in range(nsteps):
        rows = slice(s * world * B, (s + 1) * world * B)
        oracle.sgd_step(ref, x[rows].astype(np.float64), y[rows].astype(np.int64), lr)
    for n, r in zip(names, ref):
        np.testing.assert_allclose(store.read(n), r, rtol=1e-3, atol=2e-4, err_msg=f"sync {n}")
    assert store.read("global/global_step").item() == nsteps, store.read("global/global_step")
    dist.barrier()
    # same thing with the fused step kernel writing into the exchange buffer (what bench --mode sync runs)
    fs = FusedMlp2Stepper(spec, B, [store.tensor(n) for n in names], store.tensor("global/global_step"), lr, dev,
                          remote=True, grad_buffers=sg.grad_views(shapes))
    fs.cursor = torch.zeros(1, dtype=torch.int32, device=dev)
    order = torch.arange(4000, device=dev, dtype=torch.int32).view(-1, world, B)[:, rank, :].contiguous().view(-1)
    fs.cursor.fill_(nsteps)
    for s in range(nsteps, 2 * nsteps):
        fs.run(X, Y, nsteps=1, batch_idx=order)
        sg.reduce_scatter_apply(lr)
    torch.cuda.synchronize(); dist.barrier()
    for s in range(nsteps, 2 * nsteps):
        rows = slice(s * world * B, (s + 1) * world * B)
        oracle.sgd_step(ref, x[rows].astype(np.float64), y[rows].astype(np.int64), lr)
    for n, r in zip(names, ref):
        np.testing.assert_allclose(store.read(n), r, rtol=1e-3, atol=3e-4, err_msg=f"sync fused {n}")
    assert store.read("global/global_step").item() == 2 * nsteps and fs.cursor.item() == 2 * nsteps
    dist.barrier()
    # and with the exchange INSIDE the step kernel (phase D), several synchronous steps in one launch
    fx = FusedMlp2Stepper(spec, B, [store.tensor(n) for n in names], store.tensor("global/global_step"), lr, dev,
                          max_steps_per_launch=nsteps, remote=True, grad_buffers=sg.grad_views(shapes))
    fx.sync_exchange = sg.exchange_struct(lr)
    fx.cursor = torch.full((1,), 2 * nsteps, dtype=torch.int32, device=dev)
    fx.run(X, Y, nsteps=nsteps, batch_idx=order)
    torch.cuda.synchronize(); dist.barrier()
    assert sg.aborted() == 0
    for s in range(2 * nsteps, 3 * nsteps):
        rows = slice(s * world * B, (s + 1) * world * B)
        oracle.sgd_step(ref, x[rows].astype(np.float64), y[rows].astype(np.int64), lr)
    for n, r in zip(names, ref):
        np.testing.assert_allclose(store.read(n), r, rtol=1e-3, atol=4e-4, err_msg=f"sync in-kernel {n}")
    assert store.read("global/global_step").item() == 3 * nsteps
    assert fx.step[:nsteps].cpu().tolist() == list(range(2 * nsteps, 3 * nsteps))
    dist.barrier()
    sg.close()
    # the all-reduce form (what --sync runs now): every rank trains on its own copy, per-CTA flag exchange,
    # every rank sums all ranks' gradients in rank order -> all copies bit-identical and equal to the oracle
    step0 = 3 * nsteps
    for mode in ("allreduce", "allreduce_ll"):
        sa = SyncGroup(store, names, rank, world, mode=mode)
        sa.refresh_replicas()
        dist.barrier()
        fa = FusedMlp2Stepper(spec, B, sa.home_params(), sa.home("global/global_step"), lr, dev,
                              max_steps_per_launch=nsteps, remote=False, grad_buffers=sa.grad_views(shapes))
        fa.sync_exchange = sa.exchange_struct(lr)
        fa.cursor = torch.full((1,), step0, dtype=torch.int32, device=dev)
        na = 1                                            # 4000 rows = 20 synchronous steps of 2 x 100: 15 are used up
        fa.run(X, Y, nsteps=na, batch_idx=order)
        fa.run(X, Y, nsteps=na, batch_idx=order)          # a second launch continues from the first one's copies
        torch.cuda.synchronize(); dist.barrier()
        assert sa.aborted() == 0
        for s in range(step0, step0 + 2 * na):
            rows = slice(s * world * B, (s + 1) * world * B)
            oracle.sgd_step(ref, x[rows].astype(np.float64), y[rows].astype(np.int64), lr)
        mine = [p.detach().clone() for p in sa.home_params()]
        for n, r, m in zip(names, ref, mine):
            np.testing.assert_allclose(m.cpu().numpy(), r, rtol=1e-3, atol=6e-4, err_msg=f"sync {mode} {n}")
            np.testing.assert_allclose(store.read(n), r, rtol=1e-3, atol=6e-4, err_msg=f"sync {mode} shard {n}")
            other = m.clone()
            dist.broadcast(other, src=0)
            assert torch.equal(other, m), f"{mode}: replica of {n} on rank {rank} differs from rank 0's"
        step0 += 2 * na
        assert sa.home("global/global_step").item() == step0 == store.read("global/global_step").item()
        assert fa.step[:na].cpu().tolist() == [step0 - 1]
        dist.barrier()
        sa.close()
    store.close()

    # ---------------- async: conservation of pushed deltas with exact arithmetic ---------------
    model, store = build_spmd_store(spec, rank, world, dev, seed=4)
    from distributedtensorflowexample_b200.ops import cuda_ops as ops
    for n in names:                      # power-of-two test values: fp32 adds are exact, order cannot matter
        if rank == 0:
            store.tensor(n).zero_()
    dist.barrier()
    reps = 64
    for n in names:
        t = store.tensor(n).view(-1)
        g = torch.full_like(t, 0.25 * (rank + 1)) if t.device == dev else None
        g = torch.full((t.numel(),), 0.25 * (rank + 1), device=dev)
        for _ in range(reps):
            ops.push_sgd(t, g, -1.0, atomic=True)
    torch.cuda.synchronize(); dist.barrier()
    expect = -reps * 0.25 * sum(r + 1 for r in range(world))
    for n in names:
        got = store.read(n)
        assert np.all(got == expect), (n, got.min(), got.max(), expect)
    dist.barrier()

    # ---------------- async training with the fused kernel over remote shards --------------------
    if rank == 0:
        store.initialize(np.random.default_rng(1))
    dist.barrier()
    f = FusedMlp2Stepper(spec, B, [store.tensor(n) for n in names], store.tensor("global/global_step"), lr, dev,
                         max_steps_per_launch=20, remote=True)
    idx = torch.randint(0, 4000, (20 * B,), device=dev, dtype=torch.int32)
    gs0 = 0
    for it in range(3):
        f.run(X, Y, nsteps=20, batch_idx=idx)
    torch.cuda.synchronize(); dist.barrier()
    assert store.read("global/global_step").item() == 3 * 20 * world        # global_step == sum of pushes
    loss = f.loss[:20].cpu().numpy()
    assert np.isfinite(loss).all() and loss[-1] < 9.0
    steps_seen = f.step[:20].cpu().numpy()
    assert (np.diff(steps_seen) > 0).all()                                   # strictly increasing observations
    dist.barrier()
    store.close()
    # ---------------- pull fused with the forward GEMM: TMA reads a PEER shard straight into tcgen05 tiles ----
    from distributedtensorflowexample_b200.models.mlp import MLPSpec
    from distributedtensorflowexample_b200.train.gpu_step import TcStepper
    bspec = MLPSpec((784, 104, 10), param_dtype="bfloat16")
    bmodel, bstore = build_spmd_store(bspec, rank, world, dev, seed=9)
    bnames = [v.name for v in bmodel.global_trainable]
    if rank == 0:
        g = np.random.default_rng(2)
        for v in bmodel.global_trainable:
            bstore.assign(v.name, (g.normal(size=v.shape) * 0.05).astype(np.float32))
    dist.barrier()
    # with round-robin placement over 2 shards W1 lives on GPU 0 and b1 on GPU 1: each rank reads one of them remotely
    assert {bstore.shard_of(bnames[0]), bstore.shard_of(bnames[1])} == {0, 1}
    tc = TcStepper(bspec, B, [bstore.tensor(n) for n in bnames], bstore.tensor("global/global_step"), 0.0, dev)
    Xb = X[:B].bfloat16()
    tc.run(Xb, Y[:B], train=False)
    torch.cuda.synchronize()
    pn = [bstore.read(n).astype(np.float64) for n in bnames]
    logits_ref, _ = oracle.forward(pn, Xb.float().cpu().numpy().astype(np.float64))
    np.testing.assert_allclose(tc.logits.cpu().numpy(), logits_ref, rtol=3e-2, atol=3e-2)
    # and the push epilogue into the (possibly remote) bf16 shard: both ranks push the same update once
    before = [bstore.read(n) for n in bnames]
    dist.barrier()
    tc.lr = 0.5
    tc.run(Xb, Y[:B], train=True)
    torch.cuda.synchronize(); dist.barrier()
    _, _, grads = oracle.gradients(pn, Xb.float().cpu().numpy().astype(np.float64), y[:B].astype(np.int64))
    for n, b0, gr in zip(bnames, before, grads):
        want = b0 - world * 0.5 * gr
        np.testing.assert_allclose(bstore.read(n), want, rtol=0, atol=2 ** -6 * (np.abs(want).max() + 1e-3) + 0.1 * np.abs(gr).max())
    assert bstore.read("global/global_step").item() == world
    dist.barrier()
    bstore.close()
    if rank == 0:
        print("MULTIGPU_OK")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
