"""``GpuTrainer``: the public GPU training API a worker (or a benchmark) drives.

One trainer = one worker replica bound to the parameter shards.  Two ways to feed it:

* ``train_device_steps(dataset, n)`` - data already resident in HBM (``DeviceDataset``), batches are
  sampled on the device and ``n`` steps run back to back with no host involvement per step;
* ``train_host_batch(x, y)`` - the reference's feed_dict path (worker.py:133-137): one batch in
  pinned host memory is copied H2D, one step runs, and the loss comes back D2H.  Copies are
  double-buffered on a side stream so the copy of batch s+1 overlaps the compute of batch s.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

from ..models.mlp import MLPSpec
from .gpu_step import DeviceDataset, FusedMlp2Stepper, LayerwiseStepper, TcStepper


class GpuTrainer:
    def __init__(self, spec: MLPSpec, batch: int, lr: float, params: Sequence[torch.Tensor],
                 global_step: torch.Tensor, device, steps_per_launch: int = 1, atomic: bool = True,
                 engine: str = "auto", remote: bool = True, tf32: bool = False):
        self.spec, self.batch, self.lr = spec, batch, lr
        self.device = torch.device(device)
        self.params = list(params)
        self.global_step = global_step
        self.steps_per_launch = max(1, steps_per_launch)
        if engine == "auto":
            if spec.param_dtype == "bfloat16":
                engine = "tc"
            else:
                engine = "fused" if FusedMlp2Stepper.supports(spec, batch) else "layerwise"
        self.engine = engine
        self.launch_count = 0
        if engine == "fused":
            self.fused = FusedMlp2Stepper(spec, batch, self.params, global_step, lr, self.device,
                                          max_steps_per_launch=self.steps_per_launch, atomic=atomic, remote=remote)
            self.layer = LayerwiseStepper(spec, batch, self.params, global_step, lr, self.device, atomic=atomic, tf32=tf32)
        elif engine == "layerwise":
            self.fused = None
            self.layer = LayerwiseStepper(spec, batch, self.params, global_step, lr, self.device, atomic=atomic, tf32=tf32)
        elif engine == "tc":
            self.fused = None
            self.layer = TcStepper(spec, batch, self.params, global_step, lr, self.device, stage_all=remote)
        else:
            raise ValueError(engine)
        self.input_dtype = torch.bfloat16 if engine == "tc" else torch.float32
        # host-fed path state
        self._copy_stream: Optional[torch.cuda.Stream] = None
        self._stage: List[Tuple[torch.Tensor, torch.Tensor, torch.cuda.Event, torch.cuda.Event]] = []
        self._host_out: Optional[torch.Tensor] = None
        self._slot = 0

    # ---- device-resident data ---------------------------------------------------------
    def train_device_steps(self, dataset: DeviceDataset, nsteps: int) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
        """Run ``nsteps`` steps; returns device views (loss[n], acc[n], step[n]) of the LAST launch chunk
        when chunked, so callers wanting every step's metrics use nsteps <= steps_per_launch."""
        done = 0
        last = None
        while done < nsteps:
            n = min(self.steps_per_launch, nsteps - done)
            idx = dataset.next_indices(self.batch, n)
            if self.fused is not None:
                self.fused.run(dataset.images, dataset.labels, nsteps=n, batch_idx=idx)
                self.launch_count += 1
                last = (self.fused.loss[:n], self.fused.acc[:n], self.fused.step[:n])
            else:
                for s in range(n):
                    rows = idx[s * self.batch:(s + 1) * self.batch].long()
                    self.layer.run(dataset.images.index_select(0, rows), dataset.labels.index_select(0, rows))
                    self.launch_count += self.layer.launches_per_call
                last = (self.layer.loss, self.layer.acc, self.layer.step)
            done += n
        return last

    def run_indices(self, dataset: DeviceDataset, idx: torch.Tensor, nsteps: int) -> None:
        """Fused engine only: run exactly ``nsteps`` steps over precomputed batch indices."""
        assert self.fused is not None
        done = 0
        while done < nsteps:
            n = min(self.steps_per_launch, nsteps - done)
            self.fused.run(dataset.images, dataset.labels, nsteps=n,
                           batch_idx=idx[done * self.batch:(done + n) * self.batch])
            self.launch_count += 1
            done += n

    # ---- host-fed (feed_dict) path ---------------------------------------------------
    def _ensure_host_path(self, ring: int) -> None:
        if self._copy_stream is not None:
            return
        self._copy_stream = torch.cuda.Stream(device=self.device)
        D = self.spec.dims[0]
        for _ in range(2):
            self._stage.append((torch.empty(self.batch, D, device=self.device, dtype=self.input_dtype),
                                torch.empty(self.batch, device=self.device, dtype=torch.int32),
                                torch.cuda.Event(), torch.cuda.Event()))
        self._host_out = torch.zeros(ring, 3, dtype=torch.float32).pin_memory()
        self._dev_out = torch.zeros(3, device=self.device, dtype=torch.float32)

    def train_host_batch(self, x_pinned: torch.Tensor, y_pinned: torch.Tensor, out_index: int = 0,
                         ring: int = 4096) -> torch.Tensor:
        """One step fed from pinned host memory.  Returns the pinned row that will hold
        (loss, accuracy, step) once the stream has drained (call ``torch.cuda.synchronize``
        or ``finish_host_batches`` before reading)."""
        self._ensure_host_path(ring)
        xs, ys, copied, consumed = self._stage[self._slot]
        self._slot ^= 1
        main = torch.cuda.current_stream(self.device)
        with torch.cuda.stream(self._copy_stream):
            self._copy_stream.wait_event(consumed)      # the step that last used this buffer is done
            xs.copy_(x_pinned, non_blocking=True)
            ys.copy_(y_pinned, non_blocking=True)
            copied.record(self._copy_stream)
        main.wait_event(copied)
        if self.fused is not None:
            self.fused.run(xs, ys, nsteps=1, rows_per_step=0)
            self.launch_count += 1
            loss_t = self.fused.loss
        else:
            self.layer.run(xs, ys)
            self.launch_count += self.layer.launches_per_call
            loss_t = self.layer.loss
        consumed.record(main)
        row = self._host_out[out_index % self._host_out.shape[0]]
        row[0:1].copy_(loss_t[0:1], non_blocking=True)
        return row

    def train_host_batches(self, x_list, y_list, loss_host: torch.Tensor, steps_per_launch: int = 4) -> torch.Tensor:
        """Feed ``len(x_list)`` batches from pinned host memory, one training step each (the reference's
        ``sess.run(train_op, feed_dict=...)`` loop, worker.py:133-137), with the loop itself in native code
        when the fused engine is active.  ``loss_host[i]`` (pinned) holds step i's loss once the stream has
        drained.  Each batch gets its own H2D copy and each loss its own slot in the D2H read-back;
        ``steps_per_launch`` consecutive steps share one kernel launch."""
        if self.fused is not None:
            W1, b1, W2, b2 = self.params
            spl = max(1, min(16, steps_per_launch))
            self.fused.k.train_host(W1, b1, W2, b2, self.global_step, x_list, y_list, loss_host, self.lr,
                                    atomic=self.fused.atomic, remote=self.fused.remote, steps_per_launch=spl)
            self.launch_count += (len(x_list) + spl - 1) // spl
        else:
            for i, (x, y) in enumerate(zip(x_list, y_list)):
                row = self.train_host_batch(x, y, out_index=i)
                loss_host[i:i + 1].copy_(self.layer.loss, non_blocking=True)
        return loss_host

    def h2d_bytes_per_step(self) -> int:
        return self.batch * self.spec.dims[0] * (2 if self.input_dtype == torch.bfloat16 else 4) + self.batch * 4

    def d2h_bytes_per_step(self) -> int:
        return 4

    # ---- evaluation ------------------------------------------------------------------
    def evaluate(self, X: torch.Tensor, labels: torch.Tensor) -> float:
        if isinstance(self.layer, TcStepper):
            correct = total = 0
            B = self.batch
            for s in range(0, X.shape[0] - B + 1, B):           # whole batches (buffers are batch-sized)
                self.layer.run(X[s:s + B].to(torch.bfloat16), labels[s:s + B], train=False)
                correct += int(round(self.layer.acc.item() * B)); total += B
            return correct / max(total, 1)
        correct, total = self.layer.eval_counts(X, labels)
        return float(correct.item()) / float(total)

    def snapshot_params(self) -> List[np.ndarray]:
        return [(torch.cat(list(p), 0) if isinstance(p, (list, tuple)) else p).detach().float().cpu().numpy()
                for p in self.params]
