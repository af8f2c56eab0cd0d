"""Cluster description: job name -> ordered task addresses (reference utils.py:10-26,
main.py:47-48; TF1 tf.train.ClusterSpec - SURVEY C3/T1)."""
from __future__ import annotations

from typing import Dict, List, Tuple

LOCALHOST = "127.0.0.1"
BASE_PORT = 12222


def cluster_spec(num_workers: int, num_ps: int, base_port: int = BASE_PORT, host: str = LOCALHOST) -> Dict[str, List[str]]:
    """``{'ps': [...], 'worker': [...]}`` with consecutive loopback ports, ps tasks first.

    Same shape and numbering as the reference helper: ps:k listens on ``base_port+k`` and
    worker:i on ``base_port+num_ps+i``.
    """
    if num_workers < 0 or num_ps < 0:
        raise ValueError("num_workers and num_ps must be non-negative")
    ports = iter(range(base_port, base_port + num_ps + num_workers))
    return {
        "ps": [f"{host}:{next(ports)}" for _ in range(num_ps)],
        "worker": [f"{host}:{next(ports)}" for _ in range(num_workers)],
    }


class ClusterSpec:
    """Immutable view of a cluster dict; task index == position in the job's list."""

    def __init__(self, spec: "Mapping[str, Iterable[str]] | ClusterSpec"):
        if isinstance(spec, ClusterSpec):
            spec = spec.as_dict()
        self._jobs: Dict[str, Tuple[str, ...]] = {}
        for job, tasks in spec.items():
            tasks = tuple(tasks)
            for t in tasks:
                host, _, port = t.rpartition(":")
                if not host or not port.isdigit():
                    raise ValueError(f"bad task address {t!r} in job {job!r} (want host:port)")
            self._jobs[str(job)] = tasks

    @property
    def jobs(self) -> List[str]:
        return sorted(self._jobs)

    def num_tasks(self, job: str) -> int:
        return len(self._job(job))

    def task_address(self, job: str, index: int) -> str:
        tasks = self._job(job)
        if not 0 <= index < len(tasks):
            raise ValueError(f"no task {index} in job {job!r} (has {len(tasks)})")
        return tasks[index]

    def task_host_port(self, job: str, index: int) -> Tuple[str, int]:
        host, _, port = self.task_address(job, index).rpartition(":")
        return host, int(port)

    def job_tasks(self, job: str) -> List[str]:
        return list(self._job(job))

    def as_dict(self) -> Dict[str, List[str]]:
        return {j: list(t) for j, t in self._jobs.items()}

    def global_rank(self, job: str, index: int) -> int:
        """Flat rank used by the collective back-ends: ps tasks first, then workers."""
        self.task_address(job, index)
        order = ["ps"] + [j for j in self.jobs if j != "ps"]
        rank = 0
        for j in order:
            if j == job:
                return rank + index
            rank += len(self._jobs.get(j, ()))
        raise ValueError(job)

    @property
    def world_size(self) -> int:
        return sum(len(t) for t in self._jobs.values())

    def _job(self, job: str) -> Tuple[str, ...]:
        if job not in self._jobs:
            raise ValueError(f"no job {job!r} in cluster (jobs: {self.jobs})")
        return self._jobs[job]

    def __eq__(self, other) -> bool:
        return isinstance(other, ClusterSpec) and self._jobs == other._jobs

    def __repr__(self) -> str:
        return f"ClusterSpec({self.as_dict()!r})"
