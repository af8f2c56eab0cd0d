"""The comparison baseline BASELINE.json names when the TF1 reference cannot run: a parameter server
written with stock PyTorch ops and NCCL collectives, same model / config / metric as bench.py's own arm.

Each rank is a worker and owns the variables that round-robin placement gives it ("ps shard" = the
owner's copy).  One step = the reference's step (worker.py:131-141) expressed with collectives:

    pull   : dist.broadcast(param, src=owner) for every variable            (sync_op)
    compute: forward / loss / backward with torch ops (cuBLAS GEMMs, autograd-free manual backward)
    push   : dist.reduce(grad, dst=owner, SUM); owner applies p -= lr * g/W   (apply_gradients on the ps)
    step   : owner of global_step adds 1

This is the synchronous NCCL equivalent (an asynchronous one-sided update cannot be expressed with
NCCL collectives); with one GPU the collectives vanish and it is simply the eager PyTorch step.
``--graph`` style capture is applied when ``DTFE_BASELINE_GRAPH=1`` to give the baseline its best case.
Nothing from this repository's kernels or engine is on this path.
"""
from __future__ import annotations

import json
import os
import time

import torch
import torch.distributed as dist


def _dims(config):
    return (784, 100, 10) if config == "mnist" else (784, 8192, 8192, 10)


class NcclPsWorker:
    def __init__(self, dims, batch, lr, rank, world, dev, dtype):
        self.dims, self.B, self.lr, self.rank, self.world, self.dev, self.dtype = dims, batch, lr, rank, world, dev, dtype
        g = torch.Generator(device=dev).manual_seed(0)
        self.params = []
        for a, b in zip(dims[:-1], dims[1:]):
            std = 1.0 if len(dims) == 3 else 1.0 / a ** 0.5
            self.params.append((torch.randn(a, b, device=dev, generator=g) * std).to(dtype))
            self.params.append(torch.zeros(b, device=dev, dtype=dtype))
        self.owner = [i % world for i in range(len(self.params))]
        self.global_step = torch.zeros(1, device=dev, dtype=torch.int32)
        self.loss = torch.zeros((), device=dev)

    def step(self, x, y):
        W = self.world
        if W > 1:
            for p, o in zip(self.params, self.owner):
                dist.broadcast(p, src=o)
        acts = [x]
        h = x
        L = len(self.params) // 2
        for l in range(L):
            z = torch.addmm(self.params[2 * l + 1], h, self.params[2 * l])
            if l < L - 1:
                h = torch.sigmoid(z)
                acts.append(h)
        logits = z.float()
        lse = torch.logsumexp(logits, dim=1)
        self.loss = (lse - logits.gather(1, y.long().unsqueeze(1)).squeeze(1)).mean()
        d = torch.softmax(logits, dim=1)
        d.scatter_add_(1, y.long().unsqueeze(1), torch.full((x.shape[0], 1), -1.0, device=self.dev))
        d = (d / x.shape[0]).to(self.dtype)
        grads = [None] * len(self.params)
        for l in reversed(range(L)):
            h_in = acts[l]
            grads[2 * l] = h_in.t() @ d
            grads[2 * l + 1] = d.sum(0)
            if l > 0:
                d = (d @ self.params[2 * l].t()) * h_in * (1 - h_in)
        for p, g, o in zip(self.params, grads, self.owner):
            if W > 1:
                dist.reduce(g, dst=o)
            if o == self.rank:
                p.add_(g, alpha=-self.lr / W)
        if self.owner[-1] == self.rank or W == 1:
            self.global_step += 1


def run(args, rank, world, dev) -> int:
    from bench import ClockSampler
    dims = _dims(args.config)
    mnist = args.config == "mnist"
    B = args.batch or (100 if mnist else 4096)
    dtype = torch.float32 if mnist else torch.bfloat16
    K, Wm = args.steps, max(args.warmup, 3)
    wk = NcclPsWorker(dims, B, 0.001, rank, world, dev, dtype)
    gen = torch.Generator(device=dev).manual_seed(100 + rank)
    n = 55000 if mnist else 8 * B
    data = torch.rand(n, dims[0], device=dev, generator=gen).to(dtype)
    labels = torch.randint(0, dims[-1], (n,), device=dev, generator=gen, dtype=torch.int32)
    idx = torch.randint(0, n - B, (K + Wm,), generator=torch.Generator().manual_seed(rank)).tolist()

    use_graph = os.environ.get("DTFE_BASELINE_GRAPH") == "1"
    xs, ys = torch.zeros(B, dims[0], device=dev, dtype=dtype), torch.zeros(B, device=dev, dtype=torch.int32)
    graph = None
    if use_graph:
        s = torch.cuda.Stream()
        with torch.cuda.stream(s):
            for _ in range(3):
                wk.step(xs, ys)
        torch.cuda.current_stream().wait_stream(s)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            wk.step(xs, ys)

    def one(i):
        o = idx[i]
        if graph is not None:
            xs.copy_(data[o:o + B]); ys.copy_(labels[o:o + B])
            graph.replay()
        else:
            wk.step(data[o:o + B], labels[o:o + B])

    for i in range(Wm):
        one(i)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    clocks = ClockSampler(dev.index or 0, period_s=0.01)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    clocks.start()
    e0.record()
    for i in range(K):
        one(Wm + i)
    e1.record()
    torch.cuda.synchronize()
    clk = clocks.stop()
    t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())

    # end to end: pinned host batches H2D every step, loss D2H every step
    xh = [data[i * B:(i + 1) * B].cpu().pin_memory() for i in range(8)]
    yh = [labels[i * B:(i + 1) * B].cpu().pin_memory() for i in range(8)]
    lossbuf = torch.zeros(1).pin_memory()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(K):
        xs.copy_(xh[i % 8], non_blocking=True); ys.copy_(yh[i % 8], non_blocking=True)
        if graph is not None:
            graph.replay()
        else:
            wk.step(xs, ys)
        lossbuf.copy_(wk.loss.reshape(1), non_blocking=True)
    torch.cuda.synchronize()
    te = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    sec = float(te.item())
    metric = ("MNIST MLP examples/sec (device-timed, max over ranks)" if mnist
              else "wide MLP 784-8192-8192-10 examples/sec (device-timed, max over ranks)")
    out = {"metric": metric, "value": world * B * K / (ms * 1e-3), "unit": "examples/sec", "n_gpus": world,
           "steps": K, "warmup": Wm, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak",
           "vs_baseline": None, "dtype": "fp32" if mnist else "bf16", "data": "synthetic", "impl": "nccl-ps",
           "config": {"model": "MNIST MLP 784-100-10" if mnist else "wide MLP 784-8192-8192-10",
                      "global_batch": B * world, "batch_per_worker": B,
                      "parallelism": f"nccl ps dp{world}: broadcast pull + reduce push, owner applies",
                      "cuda_graph": use_graph},
           "gpu_launches": 0, "clocks": clk,
           "e2e": {"value": world * B * K / sec, "unit": "examples/sec",
                   "h2d_bytes_per_step": B * dims[0] * (4 if mnist else 2) + B * 4, "d2h_bytes_per_step": 4},
           "final_loss": float(wk.loss.item())}
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0
