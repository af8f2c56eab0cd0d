"""`bench.py --config wide`: 784-8192-8192-10 sigmoid MLP, batch 4096 per worker, bf16 parameters and
activations, fp32 accumulation - the GEMM-bound push/pull configuration of BASELINE.json.  Every GEMM is
the tcgen05 kernel with its fused epilogue; weights are read from / pushed into the parameter shards."""
from __future__ import annotations

import json
import time

import numpy as np
import torch
import torch.distributed as dist

from .benchlib import _barrier, _max_over_ranks
from .models import oracle

METRIC = "wide MLP 784-8192-8192-10 examples/sec (device-timed, max over ranks)"


def run(args, rank: int, world: int, dev, ClockSampler) -> int:
    from .models.mlp import MLPSpec
    from .parallel.spmd import build_spmd_store
    from .train.trainer import GpuTrainer

    dims = (784, 8192, 8192, 10)
    spec = MLPSpec(dims, param_dtype="bfloat16")
    B = args.batch or 4096
    lr = 0.001
    K, W = args.steps, max(args.warmup, 3)
    # Big matrices are row-partitioned over all shards (1/N of W2 in every GPU's HBM): pulls and pushes
    # become a balanced all-to-all instead of every worker hammering one owner.
    model, store = build_spmd_store(spec, rank, world, dev, seed=0, ps_strategy="greedy_bytes",
                                    partition_min_bytes=(1 << 20) if world > 1 else 0)
    names = [v.name for v in model.global_trainable]
    params = [store.param_handle(n) for n in names]
    gs = store.tensor("global/global_step")
    # keep the sigmoids out of saturation so the backward GEMMs see non-trivial data
    if rank == 0:
        g = torch.Generator(device=dev).manual_seed(1)
        for n in names:
            shape = store.layout[n].shape
            if len(shape) == 2:
                store.assign(n, (torch.randn(shape, device=dev, generator=g) * (1.0 / shape[0] ** 0.5)).cpu().numpy())
    _barrier(world)
    tr = GpuTrainer(spec, B, lr, params, gs, dev, engine="tc", remote=not store.all_local())

    nbatches = 8
    gen = torch.Generator(device=dev).manual_seed(100 + rank)
    data = torch.rand(nbatches * B, dims[0], device=dev, generator=gen).to(torch.bfloat16)
    labels = torch.randint(0, dims[-1], (nbatches * B,), device=dev, generator=gen, dtype=torch.int32)

    def step(s):
        o = (s % nbatches) * B
        tr.layer.run(data[o:o + B], labels[o:o + B])

    for s in range(W):
        step(s)
    _barrier(world)
    R = max(1, min(getattr(args, "repeats", 25), 9))
    clocks = ClockSampler(dev.index or 0, period_s=0.02)
    clocks.start()
    samples = []
    for r in range(R):
        _barrier(world)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for s in range(K):
            step(s)
        e1.record()
        torch.cuda.synchronize()
        samples.append(e0.elapsed_time(e1))
    clk = clocks.stop()
    from .benchlib import _max_over_ranks_vec
    samples = sorted(_max_over_ranks_vec(samples, world, dev))
    ms = samples[len(samples) // 2]
    _barrier(world)
    value = world * B * K / (ms * 1e-3)
    flops = oracle.flops_per_step(dims, B)["step"]

    e2e = None
    if not args.no_e2e:
        Ke = args.e2e_steps or K
        xs = [data[i * B:(i + 1) * B].cpu().pin_memory() for i in range(nbatches)]
        ys = [labels[i * B:(i + 1) * B].cpu().pin_memory() for i in range(nbatches)]
        for s in range(3):
            tr.train_host_batch(xs[s % nbatches], ys[s % nbatches], out_index=s)
        _barrier(world)
        t0 = time.perf_counter()
        for s in range(Ke):
            last = tr.train_host_batch(xs[s % nbatches], ys[s % nbatches], out_index=s)
        torch.cuda.synchronize()
        sec = _max_over_ranks(time.perf_counter() - t0, world, dev)
        assert np.isfinite(float(last[0]))
        e2e = {"value": world * B * Ke / sec, "unit": "examples/sec", "steps": Ke,
               "h2d_bytes_per_step": tr.h2d_bytes_per_step(), "d2h_bytes_per_step": tr.d2h_bytes_per_step(),
               "timing": "host wall clock around GpuTrainer.train_host_batch, max over ranks"}
        _barrier(world)

    out = {
        "metric": METRIC, "value": value, "unit": "examples/sec", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
        "data": "synthetic (uniform [0,1) 784-dim inputs, random labels), random-init weights", "impl": "ours",
        "config": {"model": "wide MLP 784-8192-8192-10 sigmoid/softmax", "global_batch": B * world,
                   "batch_per_worker": B, "seq_len": None,
                   "parallelism": f"async-ps dp{world} (shards in peer HBM; matrices >= 1 MB row-partitioned over all shards, pulled once per step, pushed from the dW epilogue)",
                   "optimizer": "sgd lr=0.001", "engine": "tc (tcgen05 GEMMs, fused pull/push epilogues)",
                   "l2": "per-step working set (147 MB weights + 2x67 MB activations + 6.4 MB batch) exceeds the 126 MB L2"},
        "gpu_launches": tr.layer.launches_per_call * K, "clocks": clk,
        "tflops_per_gpu": flops / (ms / K * 1e-3) / 1e12, "final_loss": float(tr.layer.loss.item()),
    }
    out["samples"] = {"n": len(samples), "ms_per_step_min": samples[0] / K, "ms_per_step_median": ms / K,
                      "ms_per_step_max": samples[-1] / K}
    if e2e is not None:
        out["e2e"] = e2e
    if not getattr(args, "no_baseline", False):
        # same box, same run: torch ops (cuBLAS bf16 GEMMs) + one NCCL all-reduce of the flat gradient per step,
        # whole step in a CUDA graph
        import os as _os, sys as _sys
        root = _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__)))
        if root not in _sys.path:
            _sys.path.insert(0, root)
        from baseline import nccl_ps
        try:
            m = nccl_ps.measure("wide", B, K, W, rank, world, dev, graph=True, flat=True, repeats=min(R, 5), e2e=False)
            out["baseline"] = {"impl": "nccl-ps: torch/cuBLAS bf16 GEMMs + one NCCL all-reduce per step, CUDA graph",
                               **{k: m[k] for k in ("value", "ms_per_step", "ms_per_step_min", "ms_per_step_max", "samples")}}
            out["vs_baseline"] = value / m["value"]
            out["vs_baseline_denominator"] = "in-run nccl-ps baseline (graph-captured)"
        except Exception as e:
            out["baseline"] = {"error": f"{type(e).__name__}: {e}"[:300]}
        _barrier(world)
    if rank == 0:
        print(json.dumps(out), flush=True)
    store.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0
