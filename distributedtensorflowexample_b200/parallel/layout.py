"""Byte layout of the parameter shards.

Every ps task k owns one contiguous region ("shard k").  A shard starts with a 256-byte
control header and is followed by the variables that the placement function assigned to
``/job:ps/task:k``, each aligned to 256 bytes (TMA needs 16; 256 keeps every variable on its
own L2 sector group and lets the flat push kernels vectorise).

header words (u32/int32):  [0] magic  [1] ready (chief finished init/restore; reference
``ready_op`` = report_uninitialized_variables, worker.py:112-113)  [2] stop request
[3] num variables in this shard   [4..7] reserved
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Sequence

from ..variables import VariableDef

HEADER_BYTES = 256
ALIGN = 256
MAGIC = 0x44544645  # 'DTFE'
H_MAGIC, H_READY, H_STOP, H_NVARS = 0, 1, 2, 3


@dataclass(frozen=True)
class Entry:
    name: str
    shard: int
    offset: int        # bytes from shard base
    shape: tuple
    dtype: str
    nbytes: int        # bytes reserved in the shard (rows * pitch * itemsize for 2-D tensors)
    trainable: bool
    pitch: int = 0     # row pitch in elements of 2-D tensors (0 = dense)
    parts: tuple = ()  # row-partitioned variable: ((shard, offset), ...) - part p holds rows [p*R/P, (p+1)*R/P)


def _align(n: int, a: int = ALIGN) -> int:
    return (n + a - 1) // a * a


class ShardLayout:
    def __init__(self, variables: Sequence[VariableDef], num_shards: int, partition_min_bytes: int = 0):
        """``partition_min_bytes`` > 0: 2-D variables at least this large are split by rows into one part per
        shard (TF1's partitioned variables / ``min_max_variable_partitioner`` idea), so that no single ps task
        serves - or ingests the pushes for - a whole big matrix."""
        if num_shards < 1:
            raise ValueError("num_shards must be >= 1")
        self.num_shards = num_shards
        self.partition_min_bytes = partition_min_bytes
        self.entries: Dict[str, Entry] = {}
        self.order: List[str] = []
        cursor = [HEADER_BYTES] * num_shards
        for v in variables:
            if v.device.job != "ps" or v.device.task is None:
                raise ValueError(f"variable {v.name} is not placed on the ps job (device {v.device})")
            k = v.device.task
            if not 0 <= k < num_shards:
                raise ValueError(f"variable {v.name} placed on ps task {k} but there are {num_shards} shards")
            if (partition_min_bytes and num_shards > 1 and len(v.shape) == 2 and v.nbytes >= partition_min_bytes
                    and v.shape[0] % num_shards == 0):
                item = 2 if v.dtype == "bfloat16" else 4
                pitch = (v.shape[1] + 7) // 8 * 8 if v.dtype == "bfloat16" else v.shape[1]
                part_bytes = (v.shape[0] // num_shards) * pitch * item
                parts = []
                for pk in range(num_shards):
                    o = _align(cursor[pk])
                    parts.append((pk, o))
                    cursor[pk] = o + part_bytes
                self.entries[v.name] = Entry(v.name, -1, -1, tuple(v.shape), v.dtype, part_bytes * num_shards,
                                             v.trainable, pitch if pitch != v.shape[1] else 0, tuple(parts))
                self.order.append(v.name)
                continue
            off = _align(cursor[k])
            nbytes, pitch = v.nbytes, 0
            if len(v.shape) == 2 and v.dtype == "bfloat16" and v.shape[1] % 8:
                # TMA needs 16-byte row pitches: bf16 matrices whose width is not a multiple of 8
                # (the 10-wide output layer) get padded rows inside the shard
                pitch = (v.shape[1] + 7) // 8 * 8
                nbytes = v.shape[0] * pitch * 2
            self.entries[v.name] = Entry(v.name, k, off, tuple(v.shape), v.dtype, nbytes, v.trainable, pitch)
            self.order.append(v.name)
            cursor[k] = off + nbytes
        self.shard_bytes = [_align(c) for c in cursor]

    def __getitem__(self, name: str) -> Entry:
        return self.entries[name]

    def names(self) -> List[str]:
        return list(self.order)

    def shard_entries(self, k: int) -> List[Entry]:
        return [self.entries[n] for n in self.order
                if self.entries[n].shard == k or any(pk == k for pk, _ in self.entries[n].parts)]

    def describe(self) -> str:
        lines = []
        for k in range(self.num_shards):
            lines.append(f"shard {k}: {self.shard_bytes[k]} B")
            for e in self.shard_entries(k):
                lines.append(f"  @{e.offset:>10d} {e.nbytes:>10d} B {e.dtype:8s} {e.name} {list(e.shape)}")
        return "\n".join(lines)
