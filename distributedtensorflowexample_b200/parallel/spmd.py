"""SPMD bring-up (torchrun, one rank per GPU): rank r is worker r *and* owns ps shard r.

torch.distributed (NCCL) is used for plumbing only - exchanging the 64-byte CUDA IPC handles of
the shards and for barriers around timed regions; no parameter or gradient ever travels through a
collective on the product path.
"""
from __future__ import annotations

from typing import List, Optional

import numpy as np
import torch
import torch.distributed as dist

from ..models.mlp import MLPSpec
from ..train.replica import build_replicated_model
from .ps import PeerParamStore


def build_spmd_store(spec: MLPSpec, rank: int, world: int, device: torch.device, seed: Optional[int] = 0,
                     num_ps: Optional[int] = None, ps_strategy: str = "round_robin", partition_min_bytes: int = 0):
    num_ps = world if num_ps is None else num_ps
    model = build_replicated_model(spec, "worker", rank, num_ps=num_ps, ps_strategy=ps_strategy)
    store = PeerParamStore(model.global_all, num_ps, device, partition_min_bytes=partition_min_bytes)
    mine = [k for k in range(num_ps) if k % world == rank]
    store.allocate(mine)
    if world > 1:
        gathered: List[dict] = [None] * world  # type: ignore
        dist.all_gather_object(gathered, store.export_handles())
        handles = {}
        for r, d in enumerate(gathered):
            if r != rank:
                handles.update(d)
        store.open_handles(handles)
        dist.barrier()
    # chief = task 0 initialises every shard through its mapped pointers (reference: init_op runs on
    # the chief only, worker.py:107-118); the others wait for the ready flag.
    if rank == 0:
        store.initialize(np.random.default_rng(seed))
        store.mark_ready()
    if world > 1:
        dist.barrier()
    assert store.is_ready()
    return model, store
