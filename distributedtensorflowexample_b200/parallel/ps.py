"""Parameter shards resident in GPU HBM and mapped into every worker.

The reference keeps ``global/*`` variables in the ps process's host RAM and reaches them by
name over gRPC (main.py:67-68; worker.py:24-32).  Here a ps task owns a ``cudaMalloc``'d shard;
it is exported as a CUDA IPC handle at bootstrap and opened by every worker, after which
``store.tensor(name)`` is an ordinary torch view of (possibly peer) HBM and the fused kernels
read (pull) and ``red.add`` (push) through it directly - no per-step RPC.

Three ownership models share this class:
  * single process (tests, 1-GPU bench): owns every shard;
  * SPMD (torchrun, one rank per GPU): rank r owns shard r, handles exchanged with
    ``torch.distributed.all_gather_object``;
  * launcher mode (ps processes + worker processes): ps tasks own, workers open handles obtained
    from the ps's rendezvous port (``train.server``).
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from ..variables import VariableDef
from .layout import ShardLayout, HEADER_BYTES, MAGIC, H_MAGIC, H_READY, H_STOP, H_NVARS

_TORCH_DTYPES = {"float32": torch.float32, "int32": torch.int32, "bfloat16": torch.bfloat16}


class PeerParamStore:
    def __init__(self, variables: Sequence[VariableDef], num_shards: int, device: torch.device,
                 partition_min_bytes: int = 0):
        from ..ops import cuda_ops
        self._ops = cuda_ops
        self.device = torch.device(device)
        self.variables = list(variables)
        self.layout = ShardLayout(self.variables, num_shards, partition_min_bytes)
        self.num_shards = num_shards
        self._base: List[Optional[int]] = [None] * num_shards     # device pointers (local or peer-mapped)
        self._owned: List[bool] = [False] * num_shards
        self._opened: List[bool] = [False] * num_shards
        self._views: Dict[str, torch.Tensor] = {}
        self._headers: List[Optional[torch.Tensor]] = [None] * num_shards
        self._snap_stream = None

    # ---- allocation / mapping ---------------------------------------------------------
    def allocate(self, shards: Optional[Sequence[int]] = None) -> None:
        """cudaMalloc + zero the shards this process owns and stamp their headers."""
        with torch.cuda.device(self.device):
            for k in (range(self.num_shards) if shards is None else shards):
                if self._base[k] is not None:
                    raise RuntimeError(f"shard {k} already mapped")
                self._base[k] = self._ops.shard_alloc(self.layout.shard_bytes[k])
                self._owned[k] = True
                hdr = self.header(k)
                hdr[H_MAGIC] = MAGIC
                hdr[H_NVARS] = len(self.layout.shard_entries(k))
            torch.cuda.synchronize(self.device)

    def export_handles(self) -> Dict[int, bytes]:
        return {k: self._ops.ipc_export(self._base[k]) for k in range(self.num_shards) if self._owned[k]}

    def open_handles(self, handles: Dict[int, bytes]) -> None:
        with torch.cuda.device(self.device):
            for k, h in handles.items():
                if self._base[k] is not None:
                    continue
                self._base[k] = self._ops.ipc_open(h)
                self._opened[k] = True

    def map_local_pointer(self, k: int, ptr: int) -> None:
        """In-process multi-GPU (tests): another store in this process owns the shard."""
        self._base[k] = ptr

    def base_ptr(self, k: int) -> int:
        if self._base[k] is None:
            raise RuntimeError(f"shard {k} is not mapped")
        return self._base[k]

    def all_local(self) -> bool:
        """True when every shard is memory this process allocated on its own GPU."""
        return all(self._owned)

    def is_complete(self) -> bool:
        return all(b is not None for b in self._base)

    def close(self) -> None:
        self._views.clear()
        self._headers = [None] * self.num_shards
        for k in range(self.num_shards):
            if self._base[k] is None:
                continue
            try:
                if self._opened[k]:
                    self._ops.ipc_close(self._base[k])
                elif self._owned[k]:
                    self._ops.shard_free(self._base[k])
            except Exception:
                pass
            self._base[k] = None

    # ---- views ------------------------------------------------------------------------
    def header(self, k: int) -> torch.Tensor:
        if self._headers[k] is None:
            self._headers[k] = self._ops.tensor_from_ptr(self.base_ptr(k), (HEADER_BYTES // 4,), torch.int32,
                                                         self.device, owner=self)
        return self._headers[k]

    def is_partitioned(self, name: str) -> bool:
        return bool(self.layout[name].parts)

    def parts(self, name: str) -> List[torch.Tensor]:
        """Row blocks of a partitioned variable, one per shard (block p lives in shard p's HBM)."""
        key = name + "//parts"
        ts = self._views.get(key)
        if ts is None:
            e = self.layout[name]
            if not e.parts:
                raise KeyError(f"{name} is not partitioned")
            rp = e.shape[0] // len(e.parts)
            pitch = e.pitch or e.shape[1]
            ts = []
            for shard, off in e.parts:
                full = self._ops.tensor_from_ptr(self.base_ptr(shard) + off, (rp, pitch), _TORCH_DTYPES[e.dtype],
                                                 self.device, owner=self)
                ts.append(full[:, :e.shape[1]])
            self._views[key] = ts
        return ts

    def param_handle(self, name: str):
        """What a step engine binds to: a tensor, or the list of row blocks of a partitioned variable."""
        return self.parts(name) if self.is_partitioned(name) else self.tensor(name)

    def tensor(self, name: str) -> torch.Tensor:
        t = self._views.get(name)
        if t is None:
            e = self.layout[name]
            if e.parts:
                raise KeyError(f"{name} is row-partitioned over {len(e.parts)} shards; use parts()/param_handle()")
            if e.pitch:
                full = self._ops.tensor_from_ptr(self.base_ptr(e.shard) + e.offset, (e.shape[0], e.pitch),
                                                 _TORCH_DTYPES[e.dtype], self.device, owner=self)
                t = full[:, :e.shape[1]]
            else:
                t = self._ops.tensor_from_ptr(self.base_ptr(e.shard) + e.offset, e.shape if e.shape else (1,),
                                              _TORCH_DTYPES[e.dtype], self.device, owner=self)
            self._views[name] = t
        return t

    def shard_of(self, name: str) -> int:
        """Owning shard (-1 for a row-partitioned variable, which every shard owns a block of)."""
        return self.layout[name].shard

    # ---- chief duties -------------------------------------------------------------------
    def initialize(self, rng: np.random.Generator, only_owned: bool = False) -> None:
        """Run every variable's initializer (reference ``global_init_op``, worker.py:98-99)."""
        for v in self.variables:
            if only_owned and self.layout[v.name].shard >= 0 and not self._owned[self.layout[v.name].shard]:
                continue
            val = v.initializer(v.shape, v.dtype, rng)
            self.assign(v.name, val)
        torch.cuda.synchronize(self.device)

    def assign(self, name: str, value) -> None:
        if self.is_partitioned(name):
            parts = self.parts(name)
            src = torch.as_tensor(np.asarray(value)).to(device=self.device, dtype=parts[0].dtype)
            rp = parts[0].shape[0]
            for p, t in enumerate(parts):
                t.copy_(src[p * rp:(p + 1) * rp])
            return
        t = self.tensor(name)
        src = torch.as_tensor(np.asarray(value)).to(device=self.device, dtype=t.dtype)
        t.copy_(src.reshape(t.shape))

    def read(self, name: str) -> np.ndarray:
        e = self.layout[name]
        if e.parts:
            t = torch.cat([p.detach() for p in self.parts(name)], dim=0)
        else:
            t = self.tensor(name).detach()
        arr = t.float().cpu().numpy() if t.dtype == torch.bfloat16 else t.cpu().numpy()
        return arr.reshape(e.shape)

    def snapshot(self, names: Optional[Sequence[str]] = None) -> Dict[str, np.ndarray]:
        """Fuzzy snapshot of the shards for the checkpoint thread (reference 3.5: not coordinated
        with training, tensors may be from slightly different global steps).  The D2H copies run on a
        side stream so that a save never waits for - or stalls - the training stream."""
        if self._snap_stream is None:
            self._snap_stream = torch.cuda.Stream(device=self.device)
        with torch.cuda.stream(self._snap_stream):
            out = {n: self.read(n) for n in (names or self.layout.names())}
        return out

    def mark_ready(self) -> None:
        torch.cuda.synchronize(self.device)
        for k in range(self.num_shards):
            self.header(k)[H_READY] = 1
        torch.cuda.synchronize(self.device)

    def is_ready(self) -> bool:
        return all(int(self.header(k)[H_READY].item()) == 1 for k in range(self.num_shards))

    def request_stop(self) -> None:
        for k in range(self.num_shards):
            self.header(k)[H_STOP] = 1

    def clear_stop(self) -> None:
        """A new generation of workers starts against shards that outlived the previous one (the reference's
        recovery model: restart the workers, the ps stays up): forget the old generation's stop request."""
        for k in range(self.num_shards):
            self.header(k)[H_STOP] = 0
        torch.cuda.synchronize(self.device)

    def stop_requested(self) -> bool:
        return any(int(self.header(k)[H_STOP].item()) != 0 for k in range(self.num_shards))
