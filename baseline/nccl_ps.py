"""The comparison baseline BASELINE.json names when the TF1 reference cannot run: data-parallel
training of the same model written with stock PyTorch ops and NCCL collectives - same model / config /
metric as bench.py's own arm, and given every advantage a library-only implementation has:

* ONE flat fp32 (bf16 for the wide config) parameter buffer and ONE flat gradient buffer; the variables
  are views, the backward GEMMs write straight into the gradient views;
* one collective per step: ``all_reduce(SUM)`` of the flat gradient, after which every rank applies
  ``p -= lr/W * g`` to its replica (the pull of the next step is then free - the strongest form of a
  synchronous parameter server; an asynchronous one-sided update cannot be written with NCCL collectives);
* the whole step - gather of the batch, forward, loss, backward, all-reduce, apply, global_step - captured
  in ONE CUDA graph and replayed (``graph=True``); the eager form is reported next to it.

The older per-variable form (broadcast pull + reduce push to the owning rank) is kept as
``flat=False``.  Nothing from this repository's kernels or engine is on this path.
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
    def __init__(self, dims, batch, lr, rank, world, dev, dtype, flat: bool = True):
        self.dims, self.B, self.lr, self.rank, self.world, self.dev, self.dtype = dims, batch, lr, rank, world, dev, dtype
        self.flat = flat
        g = torch.Generator(device=dev).manual_seed(0)
        shapes = []
        for a, b in zip(dims[:-1], dims[1:]):
            shapes += [(a, b), (b,)]
        sizes = [(int(torch.Size(s).numel()) + 7) // 8 * 8 for s in shapes]       # 16-byte starts for bf16 too
        self.flat_p = torch.zeros(sum(sizes), device=dev, dtype=dtype)
        self.flat_g = torch.zeros(sum(sizes), device=dev, dtype=dtype)
        self.params, self.grads = [], []
        off = 0
        for s, n in zip(shapes, sizes):
            numel = int(torch.Size(s).numel())
            self.params.append(self.flat_p[off:off + numel].view(*s))
            self.grads.append(self.flat_g[off:off + numel].view(*s))
            off += n
        for l, (a, b) in enumerate(zip(dims[:-1], dims[1:])):
            std = 1.0 if len(dims) == 3 else 1.0 / a ** 0.5
            self.params[2 * l].copy_((torch.randn(a, b, device=dev, generator=g) * std).to(dtype))
        self.owner = [i % world for i in range(len(self.params))]
        self.global_step = torch.zeros(1, device=dev, dtype=torch.int32)
        self.loss = torch.zeros((), device=dev)

    def step(self, x, y):
        W = self.world
        if W > 1 and not self.flat:
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
        for l in reversed(range(L)):
            h_in = acts[l]
            torch.mm(h_in.t(), d, out=self.grads[2 * l])
            torch.sum(d, 0, out=self.grads[2 * l + 1])
            if l > 0:
                d = (d @ self.params[2 * l].t()) * h_in * (1 - h_in)
        if self.flat:
            if W > 1:
                dist.all_reduce(self.flat_g)
            self.flat_p.add_(self.flat_g, alpha=-self.lr / W)
            self.global_step += 1
            return
        for p, g, o in zip(self.params, self.grads, self.owner):
            if W > 1:
                dist.reduce(g, dst=o)
            if o == self.rank:
                p.add_(g, alpha=-self.lr / W)
        if self.owner[-1] == self.rank or W == 1:
            self.global_step += 1


def measure(config: str, batch: int, K: int, Wm: int, rank: int, world: int, dev, graph: bool = True,
            flat: bool = True, repeats: int = 1, e2e: bool = True, ClockSampler=None) -> dict:
    """Device-timed K steps (max over ranks, median of ``repeats`` samples) and the host-fed end-to-end rate of
    the NCCL baseline.  Collective: every rank must call it with the same arguments."""
    dims = _dims(config)
    mnist = config == "mnist"
    B = batch or (100 if mnist else 4096)
    dtype = torch.float32 if mnist else torch.bfloat16
    wk = NcclPsWorker(dims, B, 0.001, rank, world, dev, dtype, flat=flat)
    gen = torch.Generator(device=dev).manual_seed(100 + rank)
    n = 55000 if mnist else 8 * B
    data = torch.rand(n, dims[0], device=dev, generator=gen).to(dtype)
    labels = torch.randint(0, dims[-1], (n,), device=dev, generator=gen, dtype=torch.int32)
    # batch rows are chosen on the device (a cursor into a precomputed offset table), so the captured graph
    # needs no host-side copy per step
    offs = torch.randint(0, n - B, (4096,), generator=torch.Generator().manual_seed(rank)).to(dev)
    cursor = torch.zeros(1, device=dev, dtype=torch.long)
    ar = torch.arange(B, device=dev)
    xs, ys = torch.zeros(B, dims[0], device=dev, dtype=dtype), torch.zeros(B, device=dev, dtype=torch.int32)

    def fetch_and_step():
        rows = offs[cursor % 4096] + ar
        torch.index_select(data, 0, rows, out=xs)
        torch.index_select(labels, 0, rows, out=ys)
        cursor.add_(1)
        wk.step(xs, ys)

    g = None
    if graph:
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(3):
                fetch_and_step()
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            fetch_and_step()
    one = g.replay if g is not None else fetch_and_step

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(Wm):
        one()
    clocks = ClockSampler(dev.index or 0, period_s=0.01) if ClockSampler is not None else None
    if clocks:
        clocks.start()
    samples = []
    for _ in range(max(1, repeats)):
        sync_all()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(K):
            one()
        e1.record()
        torch.cuda.synchronize()
        samples.append(e0.elapsed_time(e1))
    clk = clocks.stop() if clocks else None
    t = torch.tensor(samples, device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_all = sorted(t.tolist())
    ms = ms_all[len(ms_all) // 2]
    out = {"value": world * B * K / (ms * 1e-3), "unit": "examples/sec", "ms_per_step": ms / K, "cuda_graph": graph,
           "flat_allreduce": flat, "samples": len(ms_all), "ms_per_step_min": ms_all[0] / K,
           "ms_per_step_max": ms_all[-1] / K, "batch_per_worker": B, "dims": list(dims), "clocks": clk}
    if e2e:
        # end to end: pinned host batches H2D every step, loss D2H every step
        xh = [data[i * B:(i + 1) * B].cpu().pin_memory() for i in range(8)]
        yh = [labels[i * B:(i + 1) * B].cpu().pin_memory() for i in range(8)]
        lossbuf = torch.zeros(max(K, 1)).pin_memory()
        ge = None
        if graph:                              # the step alone (its inputs arrive from the host)
            ge = torch.cuda.CUDAGraph()
            with torch.cuda.graph(ge):
                wk.step(xs, ys)
        sync_all()
        t0 = time.perf_counter()
        for i in range(K):
            xs.copy_(xh[i % 8], non_blocking=True); ys.copy_(yh[i % 8], non_blocking=True)
            if ge is not None:
                ge.replay()
            else:
                wk.step(xs, ys)
            lossbuf[i:i + 1].copy_(wk.loss.reshape(1), non_blocking=True)
        torch.cuda.synchronize()
        te = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        out["e2e"] = {"value": world * B * K / float(te.item()), "unit": "examples/sec",
                      "h2d_bytes_per_step": B * dims[0] * (4 if mnist else 2) + B * 4, "d2h_bytes_per_step": 4}
    out["final_loss"] = float(wk.loss.item())
    return out


def run(args, rank, world, dev) -> int:
    from bench import ClockSampler
    mnist = args.config == "mnist"
    use_graph = os.environ.get("DTFE_BASELINE_GRAPH", "1") == "1"
    flat = os.environ.get("DTFE_BASELINE_FLAT", "1") == "1"
    K, Wm = args.steps, max(args.warmup, 3)
    m = measure(args.config, args.batch, K, Wm, rank, world, dev, graph=use_graph, flat=flat, repeats=5,
                ClockSampler=ClockSampler)
    metric = ("MNIST MLP examples/sec (device-timed, max over ranks)" if mnist
              else "wide MLP 784-8192-8192-10 examples/sec (device-timed, max over ranks)")
    out = {"metric": metric, "value": m["value"], "unit": "examples/sec", "n_gpus": world,
           "steps": K, "warmup": Wm, "ms_per_step": m["ms_per_step"], "higher_is_better": True, "scaling": "weak",
           "vs_baseline": None, "dtype": "fp32" if mnist else "bf16", "data": "synthetic", "impl": "nccl-ps",
           "config": {"model": "MNIST MLP 784-100-10" if mnist else "wide MLP 784-8192-8192-10",
                      "global_batch": m["batch_per_worker"] * world, "batch_per_worker": m["batch_per_worker"],
                      "parallelism": (f"nccl dp{world}: one all-reduce of the flat gradient, every rank applies" if flat
                                      else f"nccl ps dp{world}: broadcast pull + reduce push, owner applies"),
                      "cuda_graph": use_graph},
           "gpu_launches": 0, "clocks": m["clocks"], "e2e": m.get("e2e"), "final_loss": m["final_loss"],
           "samples": {"n": m["samples"], "ms_per_step_min": m["ms_per_step_min"], "ms_per_step_max": m["ms_per_step_max"]}}
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0
