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


class LaunchTicket:
    __slots__ = ("slot", "nsteps", "event")

    def __init__(self, slot: int, nsteps: int, event):
        self.slot, self.nsteps, self.event = slot, nsteps, event


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
        if engine in ("fused", "fused_tc"):
            self.fused = FusedMlp2Stepper(spec, batch, self.params, global_step, lr, self.device,
                                          max_steps_per_launch=self.steps_per_launch, atomic=atomic, remote=remote,
                                          tc=engine == "fused_tc")
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
        self._ring_dev = None

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

    # ---- pipelined launches (what Worker.learn drives) --------------------------------------
    def launch_async(self, dataset: DeviceDataset, nsteps: int, ring_depth: int = 3, ring_cap: int = 512) -> "LaunchTicket":
        """Enqueue ``nsteps`` (<= ``ring_cap``) training steps as ONE launch group plus the device->host copy of
        their per-step (loss, accuracy, observed global step), and return without waiting.  ``collect(ticket)``
        waits for that group only, so a caller that keeps two tickets in flight never lets the GPU idle while the
        host logs, writes summaries or checks its stop conditions.  At most ``ring_depth`` tickets may be
        outstanding (their results live in a ring of that many slots)."""
        if self._ring_dev is None:
            self._ring_cap, self._ring_depth = ring_cap, ring_depth
            self._ring_dev = torch.zeros(ring_depth, 3, ring_cap, dtype=torch.float32, device=self.device)
            self._ring_host = torch.zeros(ring_depth, 3, ring_cap, dtype=torch.float32).pin_memory()
            self._ring_np = self._ring_host.numpy()
            self._ring_events = [torch.cuda.Event() for _ in range(ring_depth)]
            self._ring_next = 0
        if not 1 <= nsteps <= self._ring_cap:
            raise ValueError(f"nsteps must be in [1, {self._ring_cap}]")
        slot = self._ring_next
        self._ring_next = (slot + 1) % self._ring_depth
        dev = self._ring_dev[slot]
        lo, ao, so = dev[0], dev[1], dev[2].view(torch.int32)
        idx = dataset.next_indices(self.batch, nsteps)
        if self.fused is not None:
            self.fused.run(dataset.images, dataset.labels, nsteps=nsteps, batch_idx=idx, outs=(lo, ao, so))
            self.launch_count += 1
        else:
            for i in range(nsteps):
                rows = idx[i * self.batch:(i + 1) * self.batch].long()
                self.layer.run(dataset.images.index_select(0, rows), dataset.labels.index_select(0, rows))
                lo[i:i + 1].copy_(self.layer.loss); ao[i:i + 1].copy_(self.layer.acc); so[i:i + 1].copy_(self.layer.step)
                self.launch_count += self.layer.launches_per_call
        self._ring_host[slot].copy_(dev, non_blocking=True)
        ev = self._ring_events[slot]
        ev.record(torch.cuda.current_stream(self.device))
        return LaunchTicket(slot, nsteps, ev)

    def collect(self, ticket: "LaunchTicket"):
        """Wait for one launch group; returns numpy views (valid until ``ring_depth`` more launches) of
        loss[n] float32, accuracy[n] float32, step[n] int32."""
        ticket.event.synchronize()
        r = self._ring_np[ticket.slot]
        n = ticket.nsteps
        return r[0, :n], r[1, :n], r[2, :n].view(np.int32)

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
            # buffers are batch-sized: whole batches, then the tail rows in a zero-padded batch whose padding rows
            # are not counted (reference worker.py:152-153 evaluates ALL test images)
            B, N = self.batch, X.shape[0]
            correct = torch.zeros((), device=X.device, dtype=torch.int64)
            for s in range(0, N, B):
                m = min(B, N - s)
                xb, yb = X[s:s + m].to(torch.bfloat16), labels[s:s + m]
                if m < B:
                    xp = torch.zeros(B, X.shape[1], device=X.device, dtype=torch.bfloat16); xp[:m] = xb
                    yp = torch.zeros(B, device=X.device, dtype=labels.dtype); yp[:m] = yb
                    xb, yb = xp, yp
                self.layer.run(xb, yb, train=False)
                correct += self.layer.rowcorrect[:m].sum()
            return float(correct.item()) / max(N, 1)
        correct, total = self.layer.eval_counts(X, labels)
        return float(correct.item()) / float(total)

    def snapshot_params(self) -> List[np.ndarray]:
        return [(torch.cat(list(p), 0) if isinstance(p, (list, tuple)) else p).detach().float().cpu().numpy()
                for p in self.params]
