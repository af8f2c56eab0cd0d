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


def _median(xs):
    s = sorted(xs)
    return s[len(s) // 2]


def _max_over_ranks_vec(values, world: int, dev):
    if world == 1:
        return list(values)
    t = torch.tensor(list(values), device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.tolist()


class _HostWait:
    """Rank 0 does something alone (the launcher-path measurement starts its own processes on every GPU) while
    the other ranks wait ON THE HOST: an NCCL barrier would park a spinning kernel on their GPUs, and the
    workers' cooperative kernels need every SM."""

    def __init__(self, world: int, key: str):
        self.world, self.key = world, key
        self.store = dist.distributed_c10d._get_default_store() if world > 1 else None

    def release(self, payload: str = "1") -> None:
        if self.store is not None:
            self.store.set(self.key, payload)

    def wait(self) -> str:
        if self.store is None:
            return ""
        self.store.wait([self.key])
        return self.store.get(self.key).decode()


def visible_gpu_ids(n: int):
    """Physical ids of the first n GPUs this process may use (the launchers' GPU_ID array)."""
    import os
    vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
    ids = [p.strip() for p in vis.split(",") if p.strip()] if vis else [str(i) for i in range(torch.cuda.device_count())]
    return ids[:n]


def measure_entrypoint(num_workers: int, num_gpus: int, training_steps: int, batch: int = 100, extra=(),
                       timeout: float = 420.0, sync: bool = False) -> dict:
    """The product's own entry point: ``main.py`` started by the headless launcher (1 ps + ``num_workers`` worker
    processes, worker i on GPU_ID[i % num_gpus]) with default flags, synthetic data of MNIST's training-set shape.
    Throughput = batch x global steps/sec as the workers themselves print it (reference worker.py:143-148), median
    over all ``speed:`` lines of the second half of the run."""
    import re
    import socket
    import tempfile
    import time as _time
    from .launch import launcher
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    logdir = tempfile.mkdtemp(prefix="dtfe_entry_")
    args = ["--logdir", logdir + "/log", "--data", "synthetic", "--seed", "1", "--training_steps", str(training_steps),
            "--base_port", str(port if port + 40 < 65000 else 23456), "--recovery_wait_secs", "1",
            "--batch_size", str(batch), *(["--sync"] if sync else []), *extra]
    ids = visible_gpu_ids(num_gpus)
    num_ps = num_gpus if sync else 1
    tasks = launcher.task_commands(num_workers, num_gpus, ids, num_ps=num_ps, device="cuda", extra=args)
    cl = launcher.HeadlessCluster(tasks, echo=False)
    t0 = _time.time()
    cl.start()
    codes = cl.wait(timeout=timeout)
    wall = _time.time() - t0
    pat = re.compile(r"^step: (\d+)\t\| cost: ([-0-9.e+naif]+)\t\| speed: ([-0-9.e+naif]+)step/sec$")
    speeds, costs = [], []
    for name, lines in cl.lines.items():
        if not name.startswith("worker"):
            continue
        for l in lines:
            m = pat.match(l)
            if m and int(m.group(1)) >= training_steps // 2:
                speeds.append(float(m.group(3))); costs.append(float(m.group(2)))
    ok = all(codes.get(f"worker{i}") == 0 for i in range(num_workers))
    out = {"num_workers": num_workers, "num_gpus": num_gpus, "num_ps": num_ps, "training_steps": training_steps,
           "mode": "sync" if sync else "async", "ok": ok, "wall_s": round(wall, 2), "speed_lines": len(speeds),
           "flags": "defaults (steps_per_launch adaptive, log_every 100, eval_every 10000, checkpoint every 30 s, "
                    "per-step summaries written) + --data synthetic --training_steps"}
    if speeds:
        gsps = _median(speeds)
        # asynchronous: one global step = one worker's batch; synchronous: one aggregated update of every worker's batch
        per_step = batch * (num_workers if sync else 1)
        out.update({"global_steps_per_sec": gsps, "value": gsps * per_step, "unit": "examples/sec",
                    "last_cost": costs[-1] if costs else None})
    else:
        tail = {n: ls[-5:] for n, ls in cl.lines.items()}
        out.update({"value": None, "error": f"no speed lines; exit codes {codes}", "tail": tail})
    import shutil
    shutil.rmtree(logdir, ignore_errors=True)
    return out


def run_mnist(args, rank: int, world: int, dev, ClockSampler) -> int:
    from .data import synthetic_arrays
    from .models.mlp import reference_mlp_spec
    from .ops import cuda_ops as ops
    from .parallel.spmd import build_spmd_store
    from .train.gpu_step import DeviceDataset
    from .train.trainer import GpuTrainer

    spec = reference_mlp_spec()
    B = args.batch or 100
    lr = 0.001
    K, W = args.steps, max(args.warmup, 3)
    spl = args.steps_per_launch or min(K, 500)
    R = max(1, getattr(args, "repeats", 25))

    model, store = build_spmd_store(spec, rank, world, dev, seed=0)
    names = [v.name for v in model.global_trainable]
    params = [store.tensor(n) for n in names]
    gs = store.tensor("global/global_step")

    # synthetic 784-dim data of MNIST's training-set shape (55000 x 784 fp32 = 172 MB > 126 MB L2);
    # every step gathers 100 random rows of it, and the L2 is flushed before every timed sample.
    x, y = synthetic_arrays(55000, seed=1234 + rank)
    ds = DeviceDataset(x, y.astype(np.int32), dev, seed=rank)
    if args.mode == "sync":
        return _run_mnist_sync(args, rank, world, dev, ClockSampler, spec, model, store, names, ds, x, y, B, lr, K, W)
    engine = args.engine if args.engine != "tc" else "layerwise"
    tr = GpuTrainer(spec, B, lr, params, gs, dev, steps_per_launch=spl, engine=engine,
                    remote=not store.all_local())

    # ------------------------------ device-timed ------------------------------------------
    # R samples of EXACTLY K steps each.  Every sample: barrier + synchronize, an L2 flush (a 256 MB fill: nothing
    # of the previous sample stays in the 126 MB L2, and the launch is queued behind it instead of on an idle
    # stream), start event, [this worker's batch sampling + the K steps], stop event, synchronize.
    flush = torch.empty(64 << 20, dtype=torch.int32, device=dev)
    def timed_steps(n):
        if tr.fused is not None:
            tr.run_indices(ds, ds.next_indices(B, n), n)        # epoch-sampler work is inside the timed region
        else:
            tr.train_device_steps(ds, n)
    timed_steps(W)
    _barrier(world)
    clocks = ClockSampler(dev.index or 0)
    launches0 = tr.launch_count
    clocks.start()
    samples = []
    for r in range(R):
        _barrier(world)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ops.l2_flush(flush)
        e0.record()
        timed_steps(K)
        e1.record()
        torch.cuda.synchronize()
        samples.append(e0.elapsed_time(e1))
    clk = clocks.stop()
    launches = (tr.launch_count - launches0) // R
    samples = sorted(_max_over_ranks_vec(samples, world, dev))
    ms = samples[len(samples) // 2]
    _barrier(world)
    value = world * B * K / (ms * 1e-3)
    del flush

    # ------------------------------ end to end (host-fed) -----------------------------------
    e2e = None
    if not args.no_e2e:
        Ke = args.e2e_steps or max(K, 1000)
        ring = 64
        xs = [torch.from_numpy(x[i * B:(i + 1) * B].copy()).pin_memory() for i in range(ring)]
        ys = [torch.from_numpy(y[i * B:(i + 1) * B].astype(np.int32)).pin_memory() for i in range(ring)]
        loss_host = torch.zeros(max(Ke, W), dtype=torch.float32).pin_memory()
        espl = max(1, min(16, getattr(args, "e2e_steps_per_launch", 8)))
        tr.train_host_batches([xs[s % ring] for s in range(W)], [ys[s % ring] for s in range(W)], loss_host,
                              steps_per_launch=espl)
        xl, yl = [xs[s % ring] for s in range(Ke)], [ys[s % ring] for s in range(Ke)]
        secs = []
        for _ in range(5):
            _barrier(world)
            t0 = time.perf_counter()
            tr.train_host_batches(xl, yl, loss_host, steps_per_launch=espl)
            torch.cuda.synchronize()
            secs.append(time.perf_counter() - t0)
        sec = _median(_max_over_ranks_vec(secs, world, dev))
        assert np.isfinite(loss_host[:Ke].numpy()).all()
        # the same loop with ONE step per kernel launch (every step: its own launch, H2D copies, D2H read)
        K1 = min(Ke, 1000)
        tr.train_host_batches(xl[:W], yl[:W], loss_host, steps_per_launch=1)
        _barrier(world)
        t0 = time.perf_counter()
        tr.train_host_batches(xl[:K1], yl[:K1], loss_host, steps_per_launch=1)
        torch.cuda.synchronize()
        sec1 = _max_over_ranks(time.perf_counter() - t0, world, dev)
        e2e = {"value": world * B * Ke / sec, "unit": "examples/sec", "steps": Ke, "samples": 5,
               "value_one_step_per_launch": world * B * K1 / sec1,
               "h2d_bytes_per_step": tr.h2d_bytes_per_step(), "d2h_bytes_per_step": tr.d2h_bytes_per_step(),
               "steps_per_launch": espl,
               "timing": "host wall clock around GpuTrainer.train_host_batches (every step: H2D copy of its batch from "
                         f"pinned memory and D2H read of its loss; {espl} consecutive steps share a kernel launch), "
                         "median of 5 runs, max over ranks"}
        _barrier(world)
    final_step = int(gs.item())

    # ------------------------------ the comparison, same box, same run -------------------------
    # BASELINE.md publishes no number (the reference prints none and cannot be installed), so the denominator is
    # the stand-in BASELINE.json names: stock torch ops + NCCL, in its strongest form (flat all-reduce, whole step
    # in one CUDA graph), measured here with the same N, batch and K.
    baseline = None
    if not getattr(args, "no_baseline", False):
        import sys as _sys, os as _os
        root = _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__)))
        if root not in _sys.path:
            _sys.path.insert(0, root)
        from baseline import nccl_ps
        baseline = {}
        for tag, graph in (("graph", True), ("eager", False)):
            try:
                m = nccl_ps.measure("mnist", B, K, W, rank, world, dev, graph=graph, flat=True, repeats=min(R, 9),
                                    e2e=not args.no_e2e)
                baseline[tag] = {k: m[k] for k in ("value", "ms_per_step", "ms_per_step_min", "ms_per_step_max", "samples")}
                if "e2e" in m:
                    baseline[tag]["e2e_value"] = m["e2e"]["value"]
            except Exception as e:                      # the comparison must not take the product's number down
                baseline[tag] = {"error": f"{type(e).__name__}: {e}"[:300]}
        _barrier(world)

    # ------------------------------ the product's entry point ----------------------------------
    entry = None
    if not getattr(args, "no_entry", False):
        gate = _HostWait(world, "dtfe_bench/entry_done")
        if rank == 0:
            try:
                steps = getattr(args, "entry_steps", 0) or 300000
                entry = measure_entrypoint(world, world, steps * world, batch=B)
                if getattr(args, "entry_multi", False) and world == 4:
                    entry["reference_topology_16w_4gpu"] = measure_entrypoint(16, 4, 400000, batch=B)
            except Exception as e:
                entry = {"value": None, "error": f"{type(e).__name__}: {e}"[:300]}
            gate.release()
        else:
            gate.wait()

    out = {
        "metric": METRIC, "value": value, "unit": "examples/sec", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "fp32", "data": "synthetic (55000x784 fp32 + int labels per worker, MNIST train shape)",
        "impl": "ours",
        "samples": {"n": len(samples), "ms_per_step_min": samples[0] / K, "ms_per_step_median": ms / K,
                    "ms_per_step_max": samples[-1] / K},
        "config": {"model": "MNIST MLP 784-100-10 sigmoid/softmax (reference worker.py:46-57)",
                   "global_batch": B * world, "batch_per_worker": B, "seq_len": None,
                   "parallelism": f"async-ps dp{world} (worker+ps shard per GPU, peer HBM)",
                   "optimizer": "sgd lr=0.001", "engine": tr.engine, "steps_per_launch": spl,
                   "l2": "L2 flushed (256 MB fill) before every timed sample; 172 MB resident dataset, random rows "
                         "every step; batch sampling is inside the timed region"},
        "gpu_launches": launches, "clocks": clk, "global_step_after": final_step,
    }
    if baseline is not None:
        out["baseline"] = {"impl": "nccl-ps (baseline/nccl_ps.py): torch ops + one NCCL all-reduce of the flat gradient "
                                   "per step, same model/batch/N/K, run inside this same process",
                           **baseline}
        g = baseline.get("graph", {}).get("value")
        if g:
            out["vs_baseline"] = value / g
            out["vs_baseline_denominator"] = "in-run nccl-ps baseline, whole step CUDA-graph-captured (its best form)"
            if e2e is not None and baseline["graph"].get("e2e_value"):
                e2e["vs_baseline_e2e"] = e2e["value"] / baseline["graph"]["e2e_value"]
    if e2e is not None:
        out["e2e"] = e2e
    if entry is not None:
        out["entry"] = entry
        if entry.get("value"):
            entry["fraction_of_device_timed"] = entry["value"] / value
    if rank == 0:
        print(json.dumps(out), flush=True)
    store.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def _run_mnist_sync(args, rank, world, dev, ClockSampler, spec, model, store, names, ds, x, y, B, lr, K, W) -> int:
    """Synchronous SGD inside the fused step kernel, any number of steps per launch (parallel/sync.py, mlp2_step.cu
    phase D).  ``--sync-mode allreduce`` (default): per-CTA flag exchange, every rank sums all ranks' gradients and
    applies them to its own copy of the parameters; ``reduce_scatter``: round 1's form (grid barrier, owners reduce
    and apply, second flag round, pull)."""
    from .ops import cuda_ops as ops
    from .parallel.sync import SyncGroup
    from .train.gpu_step import FusedMlp2Stepper
    mode = getattr(args, "sync_mode", "allreduce_ll")
    sg = SyncGroup(store, names, rank, world, mode=mode)
    shapes = {n: tuple(store.tensor(n).shape) for n in names}
    spl = args.steps_per_launch or min(K, 500)
    R = max(1, getattr(args, "repeats", 25))
    if mode.startswith("allreduce"):
        sg.refresh_replicas()
        _barrier(world)
        params, gstep, remote = sg.home_params(), sg.home("global/global_step"), False
    else:
        params, gstep, remote = [store.tensor(n) for n in names], store.tensor("global/global_step"), not store.all_local()
    st = FusedMlp2Stepper(spec, B, params, gstep, lr, dev, max_steps_per_launch=spl, remote=remote,
                          grad_buffers=sg.grad_views(shapes))
    st.sync_exchange = sg.exchange_struct(lr)       # phase D of the step kernel: exchange + apply, every step

    def run_steps(n):
        done = 0
        while done < n:
            m = min(spl, n - done)
            st.run(ds.images, ds.labels, nsteps=m, batch_idx=ds.next_indices(B, m))
            done += m
        return (n + spl - 1) // spl

    run_steps(W)
    _barrier(world)
    flush = torch.empty(64 << 20, dtype=torch.int32, device=dev)
    clocks = ClockSampler(dev.index or 0)
    clocks.start()
    samples, launches = [], 0
    for r in range(R):
        _barrier(world)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ops.l2_flush(flush)
        e0.record()
        launches = run_steps(K)
        e1.record()
        torch.cuda.synchronize()
        samples.append(e0.elapsed_time(e1))
    clk = clocks.stop()
    assert sg.aborted() == 0, "a rank abandoned the exchange"
    samples = sorted(_max_over_ranks_vec(samples, world, dev))
    ms = samples[len(samples) // 2]
    _barrier(world)
    value = world * B * K / (ms * 1e-3)
    # replicas must still agree bit for bit (all-reduce mode): compare a checksum of this rank's copy with rank 0's
    agree = None
    if mode.startswith("allreduce") and world > 1:
        chk = torch.stack([p.double().sum() for p in params] + [gstep.double().sum()])
        ref = chk.clone()
        dist.broadcast(ref, src=0)
        agree = bool(torch.equal(chk, ref))
    baseline = None
    if not getattr(args, "no_baseline", False):
        import sys as _sys, os as _os
        root = _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__)))
        if root not in _sys.path:
            _sys.path.insert(0, root)
        from baseline import nccl_ps
        try:
            m = nccl_ps.measure("mnist", B, K, W, rank, world, dev, graph=True, flat=True, repeats=min(R, 9), e2e=False)
            baseline = {k: m[k] for k in ("value", "ms_per_step", "ms_per_step_min", "ms_per_step_max", "samples")}
        except Exception as e:
            baseline = {"error": f"{type(e).__name__}: {e}"[:300]}
        _barrier(world)
    out = {"metric": METRIC, "value": value, "unit": "examples/sec", "n_gpus": world,
           "steps": K, "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak",
           "vs_baseline": None, "dtype": "fp32", "data": "synthetic", "impl": "ours",
           "samples": {"n": len(samples), "ms_per_step_min": samples[0] / K, "ms_per_step_median": ms / K,
                       "ms_per_step_max": samples[-1] / K},
           "config": {"model": "MNIST MLP 784-100-10", "global_batch": B * world, "batch_per_worker": B,
                      "parallelism": f"sync-ps dp{world}: {mode} fused with the apply over peer memory, inside the step kernel",
                      "engine": "fused step kernel, exchange as its phase D", "steps_per_launch": spl, "sync_mode": mode,
                      "l2": "L2 flushed (256 MB fill) before every timed sample; batch sampling inside the timed region"},
           "gpu_launches": launches, "clocks": clk, "replicas_bit_identical": agree,
           "global_step_after": int(gstep.item())}
    if baseline is not None:
        out["baseline"] = {"impl": "nccl-ps: torch ops + one NCCL all-reduce of the flat gradient per step (the same "
                                   "synchronous algorithm), whole step CUDA-graph-captured", **baseline}
        if baseline.get("value"):
            out["vs_baseline"] = value / baseline["value"]
    if rank == 0:
        print(json.dumps(out), flush=True)
    sg.close(); store.close()
    if world > 1:
        dist.barrier(); dist.destroy_process_group()
    return 0


def run_wide(args, rank: int, world: int, dev, ClockSampler) -> int:
    from .bench_wide import run as _run
    return _run(args, rank, world, dev, ClockSampler)
