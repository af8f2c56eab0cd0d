"""Benchmark bodies behind ``bench.py`` (kept in the package so tests can import them)."""
from __future__ import annotations

import json
import time

import numpy as np
import torch
import torch.distributed as dist

METRIC = "MNIST MLP examples/sec (device-timed, max over ranks)"


def _barrier(world):
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def _max_over_ranks(value: float, world: int, dev) -> float:
    if world == 1:
        return value
    t = torch.tensor([value], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def run_mnist(args, rank: int, world: int, dev, ClockSampler) -> int:
    from .data import synthetic_arrays
    from .models.mlp import reference_mlp_spec
    from .parallel.spmd import build_spmd_store
    from .train.gpu_step import DeviceDataset
    from .train.trainer import GpuTrainer

    spec = reference_mlp_spec()
    B = args.batch or 100
    lr = 0.001
    K, W = args.steps, max(args.warmup, 3)
    spl = args.steps_per_launch or min(K, 500)

    model, store = build_spmd_store(spec, rank, world, dev, seed=0)
    names = [v.name for v in model.global_trainable]
    params = [store.tensor(n) for n in names]
    gs = store.tensor("global/global_step")

    # synthetic 784-dim data of MNIST's training-set shape (55000 x 784 fp32 = 172 MB > 126 MB L2);
    # every step gathers 100 random rows of it, so inputs never sit in L2 from the previous step.
    x, y = synthetic_arrays(55000, seed=1234 + rank)
    ds = DeviceDataset(x, y.astype(np.int32), dev, seed=rank)
    if args.mode == "sync":
        return _run_mnist_sync(args, rank, world, dev, ClockSampler, spec, model, store, names, ds, x, y, B, lr, K, W)
    engine = args.engine if args.engine != "tc" else "layerwise"
    tr = GpuTrainer(spec, B, lr, params, gs, dev, steps_per_launch=spl, engine=engine,
                    remote=not store.all_local())

    # ------------------------------ device-timed ------------------------------------------
    if tr.fused is not None:
        idx_w = ds.next_indices(B, W)
        idx = ds.next_indices(B, K)
        tr.run_indices(ds, idx_w, W)
    else:
        tr.train_device_steps(ds, W)
    _barrier(world)
    clocks = ClockSampler(dev.index or 0)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = tr.launch_count
    clocks.start()
    e0.record()
    if tr.fused is not None:
        tr.run_indices(ds, idx, K)
    else:
        tr.train_device_steps(ds, K)
    e1.record()
    torch.cuda.synchronize()
    clk = clocks.stop()
    launches = tr.launch_count - launches0
    ms = _max_over_ranks(e0.elapsed_time(e1), world, dev)
    _barrier(world)
    value = world * B * K / (ms * 1e-3)

    # ------------------------------ end to end (host-fed) -----------------------------------
    e2e = None
    if not args.no_e2e:
        Ke = args.e2e_steps or K
        ring = 64
        xs = [torch.from_numpy(x[i * B:(i + 1) * B].copy()).pin_memory() for i in range(ring)]
        ys = [torch.from_numpy(y[i * B:(i + 1) * B].astype(np.int32)).pin_memory() for i in range(ring)]
        loss_host = torch.zeros(max(Ke, W), dtype=torch.float32).pin_memory()
        espl = max(1, min(16, getattr(args, "e2e_steps_per_launch", 8)))
        tr.train_host_batches([xs[s % ring] for s in range(W)], [ys[s % ring] for s in range(W)], loss_host,
                              steps_per_launch=espl)
        _barrier(world)
        xl, yl = [xs[s % ring] for s in range(Ke)], [ys[s % ring] for s in range(Ke)]
        t0 = time.perf_counter()
        tr.train_host_batches(xl, yl, loss_host, steps_per_launch=espl)
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        sec = _max_over_ranks(t1 - t0, world, dev)
        assert np.isfinite(loss_host[:Ke].numpy()).all()
        # the same loop with ONE step per kernel launch (every step: its own launch, H2D copies, D2H read)
        K1 = min(Ke, 1000)
        tr.train_host_batches(xl[:W], yl[:W], loss_host, steps_per_launch=1)
        _barrier(world)
        t0 = time.perf_counter()
        tr.train_host_batches(xl[:K1], yl[:K1], loss_host, steps_per_launch=1)
        torch.cuda.synchronize()
        sec1 = _max_over_ranks(time.perf_counter() - t0, world, dev)
        e2e = {"value": world * B * Ke / sec, "unit": "examples/sec", "steps": Ke,
               "value_one_step_per_launch": world * B * K1 / sec1,
               "h2d_bytes_per_step": tr.h2d_bytes_per_step(), "d2h_bytes_per_step": tr.d2h_bytes_per_step(),
               "steps_per_launch": espl,
               "timing": "host wall clock around GpuTrainer.train_host_batches (every step: H2D copy of its batch from "
                         f"pinned memory and D2H read of its loss; {espl} consecutive steps share a kernel launch), max over ranks"}
        _barrier(world)

    final_step = int(gs.item())
    out = {
        "metric": METRIC, "value": value, "unit": "examples/sec", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "fp32", "data": "synthetic (55000x784 fp32 + int labels per worker, MNIST train shape)",
        "impl": "ours",
        "config": {"model": "MNIST MLP 784-100-10 sigmoid/softmax (reference worker.py:46-57)",
                   "global_batch": B * world, "batch_per_worker": B, "seq_len": None,
                   "parallelism": f"async-ps dp{world} (worker+ps shard per GPU, peer HBM)",
                   "optimizer": "sgd lr=0.001", "engine": tr.engine, "steps_per_launch": spl,
                   "l2": "inputs larger than L2: 172 MB resident dataset, random rows every step"},
        "gpu_launches": launches, "clocks": clk, "global_step_after": final_step,
    }
    if e2e is not None:
        out["e2e"] = e2e
    if rank == 0:
        print(json.dumps(out), flush=True)
    store.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def _run_mnist_sync(args, rank, world, dev, ClockSampler, spec, model, store, names, ds, x, y, B, lr, K, W) -> int:
    """Synchronous SGD: the fused step kernel writes its gradients to this rank's peer-visible exchange buffer
    and, as its last phase, runs barrier + reduce-scatter + apply + barrier over peer memory - every step,
    any number of steps per launch (parallel/sync.py, mlp2_step.cu phase D)."""
    from .parallel.sync import SyncGroup
    from .train.gpu_step import FusedMlp2Stepper
    sg = SyncGroup(store, names, rank, world)
    shapes = {n: tuple(store.tensor(n).shape) for n in names}
    spl = args.steps_per_launch or min(K, 500)
    st = FusedMlp2Stepper(spec, B, [store.tensor(n) for n in names], store.tensor("global/global_step"), lr, dev,
                          max_steps_per_launch=spl, remote=not store.all_local(), grad_buffers=sg.grad_views(shapes))
    idx = ds.next_indices(B, K + W)
    st.sync_exchange = sg.exchange_struct(lr)       # phase D of the step kernel: exchange + apply, every step

    def run_steps(first, n):
        done = 0
        while done < n:
            m = min(spl, n - done)
            st.run(ds.images, ds.labels, nsteps=m, batch_idx=idx[(first + done) * B:(first + done + m) * B])
            done += m
        return (n + spl - 1) // spl

    run_steps(0, W)
    _barrier(world)
    clocks = ClockSampler(dev.index or 0)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    clocks.start()
    e0.record()
    launches = run_steps(W, K)
    e1.record()
    torch.cuda.synchronize()
    clk = clocks.stop()
    assert sg.aborted() == 0, "a rank abandoned the exchange"
    ms = _max_over_ranks(e0.elapsed_time(e1), world, dev)
    _barrier(world)
    out = {"metric": METRIC, "value": world * B * K / (ms * 1e-3), "unit": "examples/sec", "n_gpus": world,
           "steps": K, "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak",
           "vs_baseline": None, "dtype": "fp32", "data": "synthetic", "impl": "ours",
           "config": {"model": "MNIST MLP 784-100-10", "global_batch": B * world, "batch_per_worker": B,
                      "parallelism": f"sync-ps dp{world}: reduce-scatter fused with apply over peer memory",
                      "engine": "fused step kernel with the exchange (barrier, reduce-scatter, apply, barrier) as its phase D", "steps_per_launch": spl},
           "gpu_launches": launches, "clocks": clk,
           "global_step_after": int(store.tensor("global/global_step").item())}
    if rank == 0:
        print(json.dumps(out), flush=True)
    sg.close(); store.close()
    if world > 1:
        dist.barrier(); dist.destroy_process_group()
    return 0


def run_wide(args, rank: int, world: int, dev, ClockSampler) -> int:
    from .bench_wide import run as _run
    return _run(args, rank, world, dev, ClockSampler)
