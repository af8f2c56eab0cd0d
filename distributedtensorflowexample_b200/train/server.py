"""In-process server for one task of the cluster (reference main.py:67-68,74-75; TF1
tf.train.Server + ``.target`` - SURVEY T2, N14).

TF1 starts a gRPC master+worker service per task.  Here a task joins one of two fabrics:

* ``backend='gloo'`` (no GPU): all tasks form one torch.distributed/gloo group rendezvousing on
  ps:0's cluster-spec port; ps tasks run ``GlooParamServer.serve_forever`` in a daemon thread.
* ``backend='cuda'``: control plane = a ``TCPStore`` hosted by ps:0 on its cluster-spec port (native
  C++ store inside torch); data plane = CUDA IPC.  A ps task waits until the chief publishes the
  byte size of its shard, ``cudaMalloc``s it on its GPU, publishes the 64-byte IPC handle and then
  just keeps the memory alive.  Workers open every handle.  Nothing else ever crosses the store in
  steady state.

Like the reference's ps, the server does not block the main thread: ``main`` sleeps so that signals
stay deliverable (main.py:69-70).
"""
from __future__ import annotations

import datetime
import threading
from dataclasses import dataclass
from typing import Optional

from ..cluster import ClusterSpec


@dataclass
class ServerConfig:
    gpu_memory_fraction: Optional[float] = None    # reference main.py:64,73
    backend: str = "auto"                          # auto | cuda | gloo
    device_index: int = 0                          # "gpu:0" = the one GPU this process sees (worker.py:17-22)
    store_timeout_s: float = 600.0
    allow_soft_placement: bool = True              # reference worker.py:120: fall back to CPU when no GPU
    log_device_placement: bool = False             # reference worker.py:121 turns it off explicitly; here a flag


def resolve_backend(name: str) -> str:
    if name in ("cuda", "gloo"):
        return name
    if name in ("cpu",):
        return "gloo"
    import torch
    return "cuda" if torch.cuda.is_available() else "gloo"


class Server:
    def __init__(self, cluster, job_name: str, task_index: int, config: Optional[ServerConfig] = None,
                 start: bool = True):
        self.cluster = cluster if isinstance(cluster, ClusterSpec) else ClusterSpec(cluster)
        self.job_name, self.task_index = job_name, task_index
        self.config = config or ServerConfig()
        self.backend = resolve_backend(self.config.backend)
        self.address = self.cluster.task_address(job_name, task_index)   # validates job/task
        self.store = None
        self._thread: Optional[threading.Thread] = None
        self._shard_ptr = None
        self.error: Optional[BaseException] = None
        if start:
            self.start()

    @property
    def target(self) -> str:
        scheme = "gloo" if self.backend == "gloo" else "cuda+ipc"
        return f"{scheme}://{self.address}"

    @property
    def num_workers(self) -> int:
        return self.cluster.num_tasks("worker")

    @property
    def num_ps(self) -> int:
        return self.cluster.num_tasks("ps")

    # ------------------------------------------------------------------------------------------
    def start(self) -> None:
        if self.backend == "gloo":
            import torch
            from ..parallel import gloo_ps
            # several tasks share the host's cores (reference: many workers per machine); tiny ops
            # gain nothing from intra-op threads and oversubscription costs milliseconds per step
            torch.set_num_threads(1)
            gloo_ps.init_gloo(self.cluster, self.job_name, self.task_index, self.config.store_timeout_s)
            if self.job_name == "ps":
                srv = gloo_ps.GlooParamServer(self.task_index, self.num_workers)
                self._ps = srv
                self._thread = threading.Thread(target=self._guard(srv.serve_forever), name="ps-serve", daemon=True)
                self._thread.start()
        else:
            import torch
            from torch.distributed import TCPStore
            host, port = self.cluster.task_host_port("ps", 0)
            is_master = self.job_name == "ps" and self.task_index == 0
            self.store = TCPStore(host, port, None, is_master,
                                  timeout=datetime.timedelta(seconds=self.config.store_timeout_s),
                                  wait_for_workers=False)
            torch.cuda.set_device(self.config.device_index)
            if self.config.gpu_memory_fraction is not None:
                torch.cuda.set_per_process_memory_fraction(min(1.0, self.config.gpu_memory_fraction))
            if self.job_name == "ps":
                self._thread = threading.Thread(target=self._guard(self._host_shard), name="ps-shard", daemon=True)
                self._thread.start()

    def _guard(self, fn):
        def run():
            try:
                fn()
            except BaseException as e:   # surfaced by join()/alive()
                self.error = e
        return run

    def _host_shard(self) -> None:
        """cuda ps: allocate what the chief asks for, export it, keep it alive."""
        import torch
        from ..ops import cuda_ops
        k = self.task_index
        torch.cuda.set_device(self.config.device_index)
        nbytes = int(self.store.get(f"shard/{k}/bytes").decode())     # blocks until the chief publishes
        ptr = cuda_ops.shard_alloc(nbytes)
        torch.cuda.synchronize()
        self._shard_ptr = ptr
        self.store.set(f"shard/{k}/handle", cuda_ops.ipc_export(ptr))

    def join(self) -> None:
        """Block until the serving thread ends (gloo ps: all workers said goodbye)."""
        if self._thread is not None:
            while self._thread.is_alive():
                self._thread.join(timeout=0.5)
        if self.error is not None:
            raise self.error

    def serving(self) -> bool:
        return self._thread is not None and self._thread.is_alive()
