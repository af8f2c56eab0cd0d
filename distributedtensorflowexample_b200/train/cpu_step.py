"""CPU step engine of the gloo plumbing configuration: the reference's step, op for op, on a
*real* local replica (reference worker.py:131-141):

  sync_op   : local_v <- global_v for the 4 trainable pairs      (pull, a message per ps task)
  train_op  : loss/grads on the local replica, gradients re-paired with the global variables and
              applied ON the ps                                   (push)
  fetch     : global_step read in the same "run" as train_op (pre-increment, racy)
  counter_op: global_step += 1                                    (separate message)

Math is plain torch on CPU without autograd (the CPU stand-in for ``allow_soft_placement``).

``optimizer="adam"`` and ``sync=True`` select the ps-side modes of ``parallel/gloo_ps.py`` (not in the
reference): the worker's side of the step is unchanged except that a synchronous push blocks until the
round's mean update has been applied and returns the step, so ``counter_op`` has nothing left to do.
"""
from __future__ import annotations

import time
from typing import Dict, Tuple

import numpy as np
import torch

from ..parallel.gloo_ps import GlooParamClient
from .replica import ReplicatedModel


class CpuStepEngine:
    def __init__(self, model: ReplicatedModel, client: GlooParamClient, lr: float, push_delay_s: float = 0.0,
                 optimizer: str = "sgd", sync: bool = False, sync_timeout_s: float = 30.0,
                 first_round_grace_s: float = 0.0):
        self.model, self.client, self.lr = model, client, lr
        self.adam = {} if optimizer == "adam" else None      # default betas/eps of tf.train.AdamOptimizer
        self.sync, self.sync_timeout_s = sync, sync_timeout_s
        # The first round also has to wait for workers that are still polling for the chief's initialisation
        # (every recovery_wait_secs, 30 s by default like TF1): they join up to one poll period late.
        self._first_round_grace_s = first_round_grace_s
        self._rounds = 0
        self.pairs = model.sync_pairs()
        self.local: Dict[str, torch.Tensor] = {}
        self.push_delay_s = push_delay_s

    def local_init_op(self, rng: np.random.Generator) -> None:
        """reference ``local_init_op`` (worker.py:100): initialise the private replica."""
        for v in self.model.local_all:
            self.local[v.name] = torch.from_numpy(v.initializer(v.shape, v.dtype, rng))

    def sync_op(self) -> None:
        pulled = self.client.pull()
        for loc, glo in self.pairs:
            self.local[loc.name] = pulled[glo.name].clone()

    def _forward(self, x: torch.Tensor):
        acts = [x]
        h = x
        L = self.model.spec.num_layers
        names = [l.name for l, _ in self.pairs]
        for l in range(L):
            z = h @ self.local[names[2 * l]] + self.local[names[2 * l + 1]]
            if l < L - 1:
                h = torch.sigmoid(z)
                acts.append(h)
            else:
                return z, acts
        raise AssertionError

    def train_op(self, batch_x: np.ndarray, batch_y: np.ndarray) -> Tuple[float, float, int]:
        """Returns (cost, training-batch accuracy, observed global step)."""
        x = torch.from_numpy(np.ascontiguousarray(batch_x, dtype=np.float32))
        y = torch.from_numpy(np.ascontiguousarray(batch_y, dtype=np.float32))   # one-hot float64 -> float32, like the feed
        logits, acts = self._forward(x)
        lse = torch.logsumexp(logits, dim=1)
        loss = (lse - (y * logits).sum(1)).mean()
        probs = torch.softmax(logits, dim=1)
        acc = (probs.argmax(1) == y.argmax(1)).float().mean()
        d = (probs - y) / x.shape[0]
        names_local = [l.name for l, _ in self.pairs]
        names_global = [g.name for _, g in self.pairs]
        grads: Dict[str, torch.Tensor] = {}
        L = self.model.spec.num_layers
        for l in reversed(range(L)):
            h_in = acts[l]
            grads[names_global[2 * l]] = h_in.t() @ d
            grads[names_global[2 * l + 1]] = d.sum(0)
            if l > 0:
                d = (d @ self.local[names_local[2 * l]].t()) * h_in * (1 - h_in)
        if self.push_delay_s:
            time.sleep(self.push_delay_s)
        if self.sync:
            # the answer is the global step after this round's single mean update: the step this worker's
            # gradients were computed for is the one before it
            timeout = self.sync_timeout_s + (self._first_round_grace_s if self._rounds == 0 else 0.0)
            step = self.client.push_opt(grads, self.lr, adam=self.adam, sync=True, timeout_s=timeout) - 1
            self._rounds += 1
        else:
            if self.adam is not None:
                self.client.push_opt(grads, self.lr, adam=self.adam)
            else:
                self.client.push(grads, self.lr)
            step = self.client.read_step()
        return float(loss), float(acc), step

    def counter_op(self) -> int:
        if self.sync:
            return -1        # the ps counted the aggregated update when it applied it
        return self.client.increment_step(1)

    def accuracy(self, images: np.ndarray, labels: np.ndarray) -> float:
        x = torch.from_numpy(np.ascontiguousarray(images, dtype=np.float32))
        logits, _ = self._forward(x)
        lab = torch.from_numpy(np.asarray(labels))
        ids = lab.argmax(1) if lab.dim() == 2 else lab.long()
        return float((logits.argmax(1) == ids).float().mean())
