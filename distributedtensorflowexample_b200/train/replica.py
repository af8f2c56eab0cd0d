"""Between-graph replication builder (reference worker.py:7-43 ``Worker.__init__``).

Declares two copies of the network in one ``VariableStore``:

* scope ``global`` under ``replica_device_setter(num_ps, worker_device='/job:worker/task:i/cpu:0')``
  -> the shared copy on the ps job, plus ``global/global_step`` (int32 scalar, init 0,
  non-trainable; worker.py:29-31);
* scope ``local`` under ``/job:worker/task:i/gpu:0`` -> the worker's private replica.

On the GPU path the local replica is *virtual*: the pull (worker.py:81-85) is fused into the
kernels, which read the global copy straight from the peer-mapped shard, so no local weights are
materialised.  On the CPU/gloo path the local replica is a real set of tensors refreshed by
``sync_op`` each step.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Tuple

from ..models.mlp import MLPSpec, build_mlp_variables
from ..placement import make_ps_strategy, replica_device_setter
from ..variables import VariableDef, VariableStore, constant_initializer, get_vars


@dataclass
class ReplicatedModel:
    spec: MLPSpec
    store: VariableStore
    global_step: VariableDef
    global_device: str
    local_device: str
    num_ps: int

    @property
    def global_trainable(self) -> List[VariableDef]:
        return get_vars(self.store, "global")

    @property
    def local_trainable(self) -> List[VariableDef]:
        return get_vars(self.store, "local")

    @property
    def global_all(self) -> List[VariableDef]:
        """The saver's var_list: 4 weights + global_step (reference worker.py:102-103)."""
        return get_vars(self.store, "global", trainable=False)

    @property
    def local_all(self) -> List[VariableDef]:
        return get_vars(self.store, "local", trainable=False)

    def sync_pairs(self) -> List[Tuple[VariableDef, VariableDef]]:
        """(local, global) pairs of ``sync_op`` / ``train_op`` - pairing by creation order, like
        the reference's ``zip(get_vars('local'), get_vars('global'))`` (worker.py:76,82-85)."""
        loc, glo = self.local_trainable, self.global_trainable
        if len(loc) != len(glo):
            raise RuntimeError("local/global replicas differ in variable count")
        for l, g in zip(loc, glo):
            if l.shape != g.shape:
                raise RuntimeError(f"pairing mismatch: {l.name}{l.shape} vs {g.name}{g.shape}")
        return list(zip(loc, glo))


ADAM_SLOTS = ("Adam", "Adam_1", "grad_slot")   # TF1 slot names for m and v, plus the push mailbox


def build_replicated_model(spec: MLPSpec, job_name: str = "worker", task_index: int = 0, num_ps: int = 1,
                           ps_strategy: str = "round_robin", optimizer: str = "sgd") -> ReplicatedModel:
    store = VariableStore()
    global_device = f"/job:{job_name}/task:{task_index}/cpu:0"
    local_device = f"/job:{job_name}/task:{task_index}/gpu:0"
    setter = replica_device_setter(num_ps, worker_device=global_device,
                                   ps_strategy=make_ps_strategy(ps_strategy, num_ps))
    with store.device(setter):
        with store.variable_scope("global"):
            build_mlp_variables(store, spec)
            gs = store.get_variable("global_step", [], "int32", constant_initializer(0), trainable=False)
    if optimizer == "adam":
        # Optimizer state lives with the shard that owns the variable (owner-side apply): TF1 names the
        # moments <var>/Adam and <var>/Adam_1; <var>/grad_slot is the mailbox workers push into.
        for v in get_vars(store, "global"):
            with store.device(v.device):
                for slot in ADAM_SLOTS:
                    store.get_variable(f"{v.name}/{slot}", v.shape, "float32", constant_initializer(0), trainable=False)
        with store.device(gs.device):
            store.get_variable("global/adam_t", [], "int32", constant_initializer(0), trainable=False)
    elif optimizer != "sgd":
        raise ValueError(f"unknown optimizer {optimizer!r}")
    with store.device(local_device):
        with store.variable_scope("local"):
            build_mlp_variables(store, spec)
    return ReplicatedModel(spec, store, gs, global_device, local_device, num_ps)
