"""Device strings and ``replica_device_setter``-style variable placement
(reference worker.py:14,22,24-25; TF1 tf.device / tf.train.replica_device_setter - SURVEY T3/T4).

A placement function maps *what is being created* to *where it lives*: variables go to the
ps job (round-robin over ps tasks in creation order, or greedily by bytes), everything else
to the worker's own device.  In this framework "lives on /job:ps/task:k" means "is a slice of
HBM shard k", and that shard is peer-mapped into every worker.
"""
from __future__ import annotations

import re
from dataclasses import dataclass, replace
from typing import Callable, List, Optional, Sequence

_DEV_RE = re.compile(r"^(?:/job:(?P<job>[^/]+))?(?:/replica:(?P<replica>\d+))?(?:/task:(?P<task>\d+))?"
                     r"(?:/(?:device:)?(?P<dtype>[a-zA-Z]+):(?P<dindex>\d+|\*))?$")


@dataclass(frozen=True)
class DeviceSpec:
    job: Optional[str] = None
    task: Optional[int] = None
    device_type: Optional[str] = None   # 'cpu' | 'gpu'
    device_index: Optional[int] = None

    @staticmethod
    def from_string(text: str) -> "DeviceSpec":
        m = _DEV_RE.match(text or "")
        if m is None:
            raise ValueError(f"malformed device string {text!r}")
        d = m.groupdict()
        return DeviceSpec(
            job=d["job"],
            task=int(d["task"]) if d["task"] is not None else None,
            device_type=d["dtype"].lower() if d["dtype"] else None,
            device_index=int(d["dindex"]) if d["dindex"] not in (None, "*") else None,
        )

    def to_string(self) -> str:
        out = ""
        if self.job is not None:
            out += f"/job:{self.job}"
        if self.task is not None:
            out += f"/task:{self.task}"
        if self.device_type is not None:
            out += f"/{self.device_type}:{self.device_index if self.device_index is not None else 0}"
        return out

    def merge_from(self, other: "DeviceSpec") -> "DeviceSpec":
        """Fields set in ``other`` override fields of ``self`` (inner tf.device scopes win)."""
        return DeviceSpec(
            job=other.job if other.job is not None else self.job,
            task=other.task if other.task is not None else self.task,
            device_type=other.device_type if other.device_type is not None else self.device_type,
            device_index=other.device_index if other.device_index is not None else self.device_index,
        )

    def __str__(self) -> str:
        return self.to_string()


@dataclass(frozen=True)
class OpInfo:
    """What a placement function gets to see about the thing being created."""
    name: str
    is_variable: bool
    nbytes: int = 0


class RoundRobinStrategy:
    """TF1's default ps strategy: task (k mod ps_tasks) for the k-th variable created."""

    def __init__(self, num_tasks: int):
        if num_tasks < 1:
            raise ValueError("num_tasks must be >= 1")
        self._n, self._next = num_tasks, 0

    def __call__(self, op: OpInfo) -> int:
        t = self._next
        self._next = (self._next + 1) % self._n
        return t


class GreedyLoadBalancingStrategy:
    """Place each variable on the ps task with the fewest bytes so far (TF1's byte_size_load_fn variant)."""

    def __init__(self, num_tasks: int):
        if num_tasks < 1:
            raise ValueError("num_tasks must be >= 1")
        self._loads = [0] * num_tasks

    def __call__(self, op: OpInfo) -> int:
        t = min(range(len(self._loads)), key=lambda i: (self._loads[i], i))
        self._loads[t] += max(op.nbytes, 1)
        return t


def make_ps_strategy(name: str, num_tasks: int):
    if name in ("round_robin", "rr"):
        return RoundRobinStrategy(num_tasks)
    if name in ("greedy_bytes", "greedy"):
        return GreedyLoadBalancingStrategy(num_tasks)
    raise ValueError(f"unknown ps_strategy {name!r}")


def replica_device_setter(ps_tasks: int = 0, ps_device: str = "/job:ps", worker_device: str = "/job:worker",
                          ps_strategy: Optional[Callable[[OpInfo], int]] = None) -> Callable[[OpInfo], str]:
    """Return a device function: variables -> ``ps_device/task:k``, other ops -> ``worker_device``.

    ``ps_tasks == 0`` disables the ps placement (everything on ``worker_device``), as in TF1.
    """
    if ps_tasks == 0:
        return lambda op: worker_device
    strategy = ps_strategy if ps_strategy is not None else RoundRobinStrategy(ps_tasks)
    ps_spec = DeviceSpec.from_string(ps_device)
    worker_spec = DeviceSpec.from_string(worker_device)

    def device_fn(op: OpInfo) -> str:
        if op.is_variable:
            task = strategy(op)
            if not 0 <= task < ps_tasks:
                raise ValueError(f"ps_strategy returned task {task} outside [0,{ps_tasks})")
            return replace(ps_spec, task=task).to_string()
        return worker_spec.to_string()

    return device_fn


def assign_variables(var_infos: Sequence[OpInfo], device_fn: Callable[[OpInfo], str]) -> List[DeviceSpec]:
    """Run a device function over variables in creation order."""
    return [DeviceSpec.from_string(device_fn(v)) for v in var_infos]
