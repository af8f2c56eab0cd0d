"""ParamClient: the worker-side handle on the global variables (what a TF1 session gives the
reference: run(init_op), run(sync_op), run(train_op), run(counter_op), Saver reads).

Two implementations:
  * ``PeerParamClient``  - variables are slices of peer-mapped HBM shards (``parallel.ps``); pull and
    push are fused into the step kernels, so this client only does control-plane work
    (init/assign/snapshot/ready/step reads);
  * ``GlooParamClient``  - variables live in ps processes' host memory and every access is a message
    over torch.distributed/gloo (``parallel.gloo_ps``): the CPU plumbing configuration.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import numpy as np


class ParamClient:
    global_step_name = "global/global_step"

    def initialize(self, rng: np.random.Generator) -> None: raise NotImplementedError
    def assign_many(self, values: Dict[str, np.ndarray]) -> None: raise NotImplementedError
    def snapshot(self, names: Optional[Sequence[str]] = None) -> Dict[str, np.ndarray]: raise NotImplementedError
    def read_step(self) -> int: raise NotImplementedError
    def mark_ready(self) -> None: raise NotImplementedError
    def report_uninitialized_variables(self) -> List[str]: raise NotImplementedError
    def request_stop(self) -> None: ...
    def stop_requested(self) -> bool: return False
    def close(self) -> None: ...


class PeerParamClient(ParamClient):
    def __init__(self, store):
        self.store = store

    def initialize(self, rng):
        self.store.initialize(rng)

    def assign_many(self, values):
        for n, v in values.items():
            self.store.assign(n, v)

    def snapshot(self, names=None):
        return self.store.snapshot(names)

    def read_step(self) -> int:
        return int(self.store.tensor(self.global_step_name).item())

    def mark_ready(self):
        self.store.mark_ready()

    def report_uninitialized_variables(self):
        return [] if self.store.is_ready() else self.store.layout.names()

    def request_stop(self):
        self.store.request_stop()

    def stop_requested(self):
        return self.store.stop_requested()

    def close(self):
        self.store.close()
