"""Synchronous SGD over peer memory (no counterpart in the reference, which is async
only - the semantic model is TF1's SyncReplicasOptimizer: aggregate all replicas' gradients, apply the
mean once, global_step += 1 per aggregated update).

Each rank keeps a flat fp32 gradient buffer in IPC-exportable memory.  After its backward pass has
written the buffer, every rank launches ``dtfe_sync_rs_apply_f32`` once: barrier on flags in peer
memory -> each owner reduces the slices of the variables it owns out of all ranks' buffers (peer
loads over NVLink) and applies ``-lr/W * sum`` in place in its shard -> barrier.  That is a
reduce-scatter fused with the update; there is no all-gather because the next forward reads the
parameters from the owners' shards anyway (the fused pull).  One rank per GPU.

``mode="allreduce_ll"`` moves the same all-reduce over a low-latency transport: every gradient element is stored
as an 8-byte {value, step number} pair straight into every rank's receive buffer by the thread that computed it, and
consumers spin on their local buffer - no fence, no flag, no NVLink round trip on the critical path.

``mode="allreduce"`` (fused MNIST step kernel only) is the flag-based form: every rank keeps its own copy of
the parameters ("home": the shard itself where the rank owns the variable, a private replica elsewhere), each
CTA of the step kernel exchanges per-CTA flags with the same CTA of the other ranks, sums all ranks' gradients
for the elements it produced and applies them to its home copy - no grid barrier, no second flag round, no
pull.  Replicas stay bit-identical because every rank sums in rank order.
"""
from __future__ import annotations

from typing import Dict, List, Sequence

import torch
import torch.distributed as dist

from ..ops import cuda_ops as ops
from .ps import PeerParamStore


class SyncGroup:
    def __init__(self, store: PeerParamStore, trainable_names: Sequence[str], rank: int, world: int,
                 owner_of_shard=lambda k, world: k % world, exchange=None, timeout_s: float = 20.0,
                 mode: str = "reduce_scatter"):
        """``exchange(blob) -> [blob of rank 0, ..., blob of rank W-1]``; default: torch.distributed
        all_gather_object (SPMD).  Launcher-mode workers pass a TCPStore-based exchange."""
        if mode not in ("reduce_scatter", "allreduce", "allreduce_ll"):
            raise ValueError(mode)
        self.mode = mode
        self.store, self.rank, self.world = store, rank, world
        self.device = store.device
        self.names = list(trainable_names)
        # flat layout of the gradient vector: every variable starts on a 4-element boundary
        self.ranges: Dict[str, tuple] = {}
        n = 0
        for name in self.names:
            numel = store.tensor(name).numel()
            self.ranges[name] = (n, n + numel)
            n = (n + numel + 3) // 4 * 4
        self.total = n
        max_ranks = ops.lib().dtfe_sync_max_ranks()
        copies = 2 if mode == "allreduce" else 1       # all-reduce mode alternates between two gradient copies
        self._ll_ptr = 0
        self.CTA_FLAG_PITCH = 160
        with torch.cuda.device(self.device):
            self._grad_ptr = ops.shard_alloc(copies * max(self.total, 4) * 4)
            self._flag_ptr = ops.shard_alloc(2 * max_ranks * 4)
            self._cta_flag_ptr = ops.shard_alloc(max_ranks * self.CTA_FLAG_PITCH * 4) if mode == "allreduce" else 0
            if mode == "allreduce_ll":
                # receive buffer of the low-latency transport: [2 copies][world sources][total] {value, step} pairs
                self._ll_ptr = ops.shard_alloc(2 * world * max(self.total, 4) * 8)
        self.gradbuf = ops.tensor_from_ptr(self._grad_ptr, (copies * max(self.total, 4),), torch.float32, self.device,
                                           owner=self)
        self.local_sync = torch.zeros(4, dtype=torch.int32, device=self.device)
        self.epoch = 0
        # exchange IPC handles
        mine = (ops.ipc_export(self._grad_ptr), ops.ipc_export(self._flag_ptr),
                ops.ipc_export(self._cta_flag_ptr) if self._cta_flag_ptr else None,
                ops.ipc_export(self._ll_ptr) if self._ll_ptr else None)
        self._opened: List[int] = []
        self.timeout_s = timeout_s
        # workers leaving (signal, crash handler) raise the stop word of shard 0; spinning peers see it
        from .layout import H_STOP
        self._abort_ptr = store.base_ptr(0) + 4 * H_STOP
        if world > 1:
            if exchange is None:
                allh = [None] * world
                dist.all_gather_object(allh, mine)
            else:
                allh = exchange(mine)
            self.grad_ptrs, self.flag_ptrs, self.cta_flag_ptrs, self.ll_ptrs = [], [], [], []
            with torch.cuda.device(self.device):
                for r, (gh, fh, ch, lh) in enumerate(allh):
                    if r == rank:
                        self.grad_ptrs.append(self._grad_ptr); self.flag_ptrs.append(self._flag_ptr)
                        self.cta_flag_ptrs.append(self._cta_flag_ptr); self.ll_ptrs.append(self._ll_ptr)
                    else:
                        g, f = ops.ipc_open(gh), ops.ipc_open(fh)
                        self._opened += [g, f]
                        self.grad_ptrs.append(g); self.flag_ptrs.append(f)
                        if ch is not None:
                            c = ops.ipc_open(ch)
                            self._opened.append(c)
                            self.cta_flag_ptrs.append(c)
                        if lh is not None:
                            c = ops.ipc_open(lh)
                            self._opened.append(c)
                            self.ll_ptrs.append(c)
            if exchange is None:
                dist.barrier()
        else:
            self.grad_ptrs, self.flag_ptrs, self.cta_flag_ptrs = [self._grad_ptr], [self._flag_ptr], [self._cta_flag_ptr]
            self.ll_ptrs = [self._ll_ptr]
        # segments this rank owns
        self.segments = []
        for name in self.names:
            if owner_of_shard(store.shard_of(name), world) == rank:
                b, e = self.ranges[name]
                self.segments.append((store.tensor(name).view(-1), b, e))
        gs_owner = owner_of_shard(store.shard_of("global/global_step"), world)
        self.global_step = store.tensor("global/global_step") if gs_owner == rank else None
        # all-reduce mode: this rank's own copy of every variable - the shard where it owns it, a replica elsewhere
        self._home: Dict[str, torch.Tensor] = {}
        if mode in ("allreduce", "allreduce_ll"):
            for name in self.names + ["global/global_step"]:
                t = store.tensor(name)
                mine_ = owner_of_shard(store.shard_of(name), world) == rank
                self._home[name] = t if mine_ else torch.zeros(t.shape, dtype=t.dtype, device=self.device)

    def home(self, name: str) -> torch.Tensor:
        """all-reduce mode: the copy of ``name`` this rank trains on (bind the step kernel to these)."""
        return self._home[name]

    def home_params(self) -> List[torch.Tensor]:
        return [self._home[n] for n in self.names]

    def refresh_replicas(self) -> None:
        """all-reduce mode: load this rank's private replicas from the shards.  Call once the model is ready
        (after the chief's init or restore) and before the first step; in synchronous mode no shard changes
        until every rank has joined the first exchange, so the copies of all ranks agree."""
        for name, t in self._home.items():
            src = self.store.tensor(name)
            if t.data_ptr() != src.data_ptr():
                t.copy_(src)
        torch.cuda.synchronize(self.device)

    def grad_views(self, shapes: Dict[str, tuple]) -> List[torch.Tensor]:
        """Per-variable views into the flat gradient buffer, in ``names`` order."""
        out = []
        for name in self.names:
            b, e = self.ranges[name]
            out.append(self.gradbuf[b:e].view(*shapes[name]))
        return out

    def reduce_scatter_apply(self, lr: float, device_epoch: bool = True) -> None:
        """``device_epoch``: the kernel numbers the exchanges itself (local_sync[3]), so this call can be
        captured in a CUDA graph and replayed."""
        if self.mode != "reduce_scatter":
            raise RuntimeError("the stand-alone exchange kernel is the reduce-scatter form; this group is all-reduce")
        self.epoch += 1
        ops.sync_rs_apply(self.grad_ptrs, self.flag_ptrs, self.rank, self.world, self.segments,
                          -lr / self.world, 0 if device_epoch else self.epoch, self.local_sync, self.global_step,
                          abort_flag_ptr=self._abort_ptr, timeout_s=self.timeout_s)

    def exchange_struct(self, lr: float):
        """Arguments for the exchange run INSIDE the fused step kernel (``dtfe_mlp2_launch.sync_exchange``): same
        buffers, flags and epoch counter as ``reduce_scatter_apply``, so the two forms can be mixed."""
        x = ops.SyncExchange()
        for r in range(self.world):
            x.grads[r] = self.grad_ptrs[r]
            x.flags[r] = self.flag_ptrs[r]
        x.rank, x.world, x.nseg = self.rank, self.world, len(self.segments)
        for i, (t, b, e) in enumerate(self.segments):
            x.seg_param[i], x.seg_begin[i], x.seg_end[i] = t.data_ptr(), b, e
        x.alpha = -lr / self.world
        x.epoch = self.local_sync.data_ptr() + 12          # local_sync[3]
        x.status = self.local_sync.data_ptr() + 8          # local_sync[2]
        x.abort_flag = self._abort_ptr
        x.timeout_ns = int(self.timeout_s * 1e9)
        x.global_step = self.global_step.data_ptr() if self.global_step is not None else None
        if self.mode == "allreduce":
            x.mode, x.grad_stride = 1, max(self.total, 4)
            for r in range(self.world):
                x.cta_flags[r] = self.cta_flag_ptrs[r]
        elif self.mode == "allreduce_ll":
            x.mode, x.grad_stride = 2, max(self.total, 4)
            for r in range(self.world):
                x.ll_recv[r] = self.ll_ptrs[r]
        self._xs = x                                       # keep the ctypes object alive
        return x

    def aborted(self) -> int:
        """0 if the last reduce_scatter_apply completed, else which barrier gave up (needs a sync)."""
        return int(self.local_sync[2].item())

    def close(self) -> None:
        for p in self._opened:
            try:
                ops.ipc_close(p)
            except Exception:
                pass
        self._opened = []
