"""CPU plumbing configuration: 1..P ps processes + W workers over torch.distributed/gloo
(BASELINE.json config 1; reference run_single_gpu.sh with ``allow_soft_placement`` and no GPU,
worker.py:120 - SURVEY 4.2, 5.8 last row).

This reproduces the reference's *message pattern* literally: per step a worker does a pull
(4 tensors ps->worker), a push (4 gradient tensors worker->ps, applied ON the ps with
``var -= lr*grad``, no locking across workers), a read of global_step, and a separate increment
(reference worker.py:131,135-137,141).  The ps is a passive server loop, like TF1's worker service.

Wire protocol (all tensors CPU): a request is an int64[4] header {op, a0, seq, a3} sent to the ps
rank, optionally followed by a payload; the ps answers on the same pair.  Each ps task owns the
variables the placement function gave it, stored as one flat float32 vector + int32 global_step.

Beyond the reference (which has plain asynchronous SGD only), ``OP_PUSH_OPT`` carries the two optional modes
the CUDA path has, so that the plumbing configuration covers the same flags:
  * Adam: the moments are ordinary ps variables (TF1 slot names ``<var>/Adam``, ``<var>/Adam_1``, placed with
    their variable), the ps applies ``m, v, var`` updates per push and counts ``global/adam_t``;
  * synchronous SGD/Adam: the ps accumulates the pushes of all workers, applies the MEAN once, bumps
    ``global_step`` once and only then answers the waiting workers (TF1's SyncReplicasOptimizer semantics).
"""
from __future__ import annotations

import functools
import signal
import struct
import threading
import time
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch
import torch.distributed as dist

from ..cluster import ClusterSpec
from ..variables import VariableDef
from .client import ParamClient

OP_PULL, OP_PUSH, OP_STEP_READ, OP_STEP_INC, OP_ASSIGN, OP_IS_READY, OP_SET_READY, OP_SHUTDOWN, OP_STOP_REQ, \
    OP_STOP_QUERY, OP_ALLOC, OP_PUSH_OPT, OP_ADAM_T_READ = range(1, 14)

F_ADAM, F_SYNC = 1, 2            # OP_PUSH_OPT flag bits (header word 3: flags | nvars << 8)
SYNC_ABORTED = -1                # reply of a synchronous push that was released without an update


class SyncAborted(RuntimeError):
    """A synchronous round could not complete (a peer left, stop was requested, or the wait timed out)."""


_SHUTDOWN_SIGNALS = {signal.SIGHUP, signal.SIGINT, signal.SIGTERM}


def _float_to_bits(x: float) -> int:
    return struct.unpack("<q", struct.pack("<d", x))[0]


def _bits_to_float(b: int) -> float:
    return struct.unpack("<d", struct.pack("<q", b))[0]


class _TaskLayout:
    """Flat layout of the float variables a ps task owns (+ whether it owns global_step)."""

    def __init__(self, variables: Sequence[VariableDef], task: int):
        self.task = task
        self.float_vars: List[VariableDef] = []
        self.offsets: Dict[str, int] = {}
        self.has_step = False
        self.has_adam_t = False
        n = 0
        for v in variables:
            if v.device.job != "ps" or v.device.task != task:
                continue
            if v.dtype == "int32":               # counters live outside the flat float vector
                if v.name.endswith("adam_t"):
                    self.has_adam_t = True
                else:
                    self.has_step = True
                continue
            self.float_vars.append(v)
            self.offsets[v.name] = n
            n += v.numel
        self.numel = n
        self.trainable = [v for v in self.float_vars if v.trainable]

    def push_descriptor(self, adam: bool) -> List[int]:
        """Per trainable variable of this task: offset of the variable, of its two moments (-1 without Adam)
        and its element count - everything the model-agnostic ps needs to apply a compact gradient vector."""
        d: List[int] = []
        for v in self.trainable:
            m = self.offsets.get(v.name + "/Adam", -1) if adam else -1
            w = self.offsets.get(v.name + "/Adam_1", -1) if adam else -1
            if adam and (m < 0 or w < 0):
                raise ValueError(f"{v.name}: the model was built without Adam slots")
            d += [self.offsets[v.name], m, w, v.numel]
        return d


def init_gloo(cluster: ClusterSpec, job: str, task: int, timeout_s: float = 600.0) -> int:
    """Join the gloo group; rank = ps tasks first, then workers.  Rendezvous on ps:0's port."""
    import datetime
    host, port = cluster.task_host_port("ps", 0)
    rank = cluster.global_rank(job, task)
    if not dist.is_initialized():
        dist.init_process_group("gloo", init_method=f"tcp://{host}:{port}", rank=rank,
                                world_size=cluster.world_size, timeout=datetime.timedelta(seconds=timeout_s))
    return rank


class GlooParamServer:
    """The ps role: owns state, serves requests until told to shut down (reference main.py:66-70)."""

    def __init__(self, task: int, num_workers: int, variables: Optional[Sequence[VariableDef]] = None):
        # Like a TF1 ps, the server knows nothing about the model until a worker (the chief) tells it
        # how much state to hold (OP_ALLOC); passing `variables` pre-sizes it (tests).
        self.task = task
        self.numel = _TaskLayout(variables, task).numel if variables is not None else 0
        self.flat = torch.zeros(max(self.numel, 1), dtype=torch.float32)
        self.step = torch.zeros(1, dtype=torch.int64)
        self.adam_t = 0
        self.sync_sum: Optional[torch.Tensor] = None       # running sum of this round's gradient vectors
        self.sync_waiting: List[int] = []                   # ranks waiting for this round's answer
        self.ready = False
        self.stop_requested = False
        self.num_workers = num_workers
        self.requests_served = 0
        self.peer_errors = 0
        self.dead = set()
        self.reply_timeout_s = 5.0

    def _reply(self, t: torch.Tensor, dst: int, op: int) -> None:
        """Send a response without letting one vanished client wedge the server for everyone else."""
        import datetime
        try:
            dist.isend(t, dst=dst).wait(timeout=datetime.timedelta(seconds=self.reply_timeout_s))
        except RuntimeError as e:
            self.peer_errors += 1
            self.dead.add(dst)
            print(f"ps:{self.task}: dropping reply to rank {dst} (op {op}): {str(e).splitlines()[0]}", flush=True)

    def _apply(self, desc: List[int], grads: torch.Tensor, lr: float, adam: bool, hyper) -> None:
        """var -= lr*g (TF1 ApplyGradientDescent) or the TF1 ApplyAdam update, per descriptor entry."""
        if adam:
            b1, b2, eps = hyper
            self.adam_t += 1
            lr_t = lr * (1.0 - b2 ** self.adam_t) ** 0.5 / (1.0 - b1 ** self.adam_t)
        c = 0
        for i in range(0, len(desc), 4):
            vo, mo, wo, n = desc[i:i + 4]
            g = grads[c:c + n]
            c += n
            if adam:
                m, w = self.flat[mo:mo + n], self.flat[wo:wo + n]
                m.mul_(b1).add_(g, alpha=1.0 - b1)
                w.mul_(b2).addcmul_(g, g, value=1.0 - b2)
                self.flat[vo:vo + n].addcdiv_(m, w.sqrt().add_(eps), value=-lr_t)
            else:
                self.flat[vo:vo + n].add_(g, alpha=-lr)

    def _release_sync(self, value: int, op: int) -> None:
        waiting, self.sync_waiting, self.sync_sum = self.sync_waiting, [], None
        for r in waiting:
            self._reply(torch.tensor([value], dtype=torch.int64), r, op)

    def serve_forever(self) -> None:
        hdr = torch.zeros(4, dtype=torch.int64)
        done_workers = 0
        last_seq: Dict[int, int] = {}
        while True:
            hdr.zero_()
            try:
                src = dist.recv(hdr)
            except RuntimeError:
                # a worker process died (reference 5.3: a non-chief crash is harmless to the others);
                # its half-open pair surfaces here - keep serving everyone else
                self.peer_errors += 1
                time.sleep(0.01)
                continue
            op, a0, seq, a3 = int(hdr[0]), int(hdr[1]), int(hdr[2]), int(hdr[3])
            # every request carries a per-client sequence number; anything else (an all-zero header or a
            # replay from a connection that went away) is not a request
            if op == 0 or src in self.dead or seq != last_seq.get(src, 0) + 1:
                self.peer_errors += 1
                time.sleep(0.001)
                continue
            last_seq[src] = seq
            n = self.numel
            self.requests_served += 1
            if op == OP_ALLOC:
                if a0 != self.numel:
                    self.numel = a0
                    self.flat = torch.zeros(max(a0, 1), dtype=torch.float32)
            elif op == OP_PULL:
                self._reply(self.flat, src, op)
            elif op == OP_PUSH:
                buf = torch.empty(max(n, 1), dtype=torch.float32)
                dist.recv(buf, src=src)
                # ApplyGradientDescent on the ps: var -= lr * grad (TF1 T10); unsynchronised across workers
                self.flat.add_(buf, alpha=-_bits_to_float(a0))
            elif op == OP_PUSH_OPT:
                flags, nv = a3 & 0xff, a3 >> 8
                desc_t = torch.empty(max(4 * nv, 1), dtype=torch.int64)
                dist.recv(desc_t, src=src)
                desc = [int(x) for x in desc_t[:4 * nv]]
                total = sum(desc[3::4])
                payload = torch.empty(3 + total, dtype=torch.float32)
                dist.recv(payload, src=src)
                hyper = (float(payload[0]), float(payload[1]), float(payload[2]))
                grads, lr, adam = payload[3:], _bits_to_float(a0), bool(flags & F_ADAM)
                if not flags & F_SYNC:
                    self._apply(desc, grads, lr, adam, hyper)             # asynchronous: apply at once
                elif self.stop_requested:
                    self._reply(torch.tensor([SYNC_ABORTED], dtype=torch.int64), src, op)
                else:
                    self.sync_sum = grads.clone() if self.sync_sum is None else self.sync_sum.add_(grads)
                    self.sync_waiting.append(src)
                    if len(self.sync_waiting) >= self.num_workers:        # the round is complete: one mean update
                        self._apply(desc, self.sync_sum.div_(self.num_workers), lr, adam, hyper)
                        self.step += 1
                        self._release_sync(int(self.step.item()), op)
            elif op == OP_ADAM_T_READ:
                self._reply(torch.tensor([self.adam_t], dtype=torch.int64), src, op)
            elif op == OP_STEP_READ:
                self._reply(self.step, src, op)
            elif op == OP_STEP_INC:
                old = self.step.clone()
                self.step += a0
                self._reply(old, src, op)
            elif op == OP_ASSIGN:
                buf = torch.empty(max(n, 1) + 1, dtype=torch.float32)
                dist.recv(buf, src=src)
                self.flat.copy_(buf[:max(n, 1)])
                if a0 >= 0:
                    self.step[0] = a0
                if a3 >= 0:
                    self.adam_t = a3
            elif op == OP_IS_READY:
                self._reply(torch.tensor([1 if self.ready else 0], dtype=torch.int64), src, op)
            elif op == OP_SET_READY:
                self.ready = True
            elif op == OP_STOP_REQ:
                self.stop_requested = True
                self._release_sync(SYNC_ABORTED, op)      # nobody may be left waiting for a round that cannot complete
            elif op == OP_STOP_QUERY:
                self._reply(torch.tensor([1 if self.stop_requested else 0], dtype=torch.int64), src, op)
            elif op == OP_SHUTDOWN:
                done_workers += 1
                if a0 == 1 or done_workers >= self.num_workers:
                    return
            else:
                raise RuntimeError(f"ps: unknown op {op} from rank {src}")


def _locked(fn):
    # One request/response exchange at a time per client: the chief's checkpoint and step-counter
    # threads share the process group with the training thread.
    # The shutdown signals (which raise SystemExit from their handler, reference main.py:51-52) are held
    # back for the duration of an exchange so that it is never abandoned half-way, which would leave
    # the request/response streams of this worker and the ps out of step.
    @functools.wraps(fn)
    def wrapper(self, *a, **kw):
        in_main = threading.current_thread() is threading.main_thread()
        old = signal.pthread_sigmask(signal.SIG_BLOCK, _SHUTDOWN_SIGNALS) if in_main else None
        try:
            with self._lock:
                return fn(self, *a, **kw)
        finally:
            if old is not None:
                signal.pthread_sigmask(signal.SIG_SETMASK, old)
    return wrapper


class GlooParamClient(ParamClient):
    adam_t_name = "global/adam_t"

    def __init__(self, variables: Sequence[VariableDef], cluster: ClusterSpec):
        self._lock = threading.RLock()
        self.variables = [v for v in variables if v.device.job == "ps"]
        self.cluster = cluster
        self.num_ps = cluster.num_tasks("ps")
        self.layouts = [_TaskLayout(self.variables, k) for k in range(self.num_ps)]
        self.step_task = next((k for k, l in enumerate(self.layouts) if l.has_step), 0)
        self._hdr = torch.zeros(4, dtype=torch.int64)
        self._seq = [0] * self.num_ps
        self.messages = 0

    def _rank(self, k: int) -> int:
        return self.cluster.global_rank("ps", k)

    def _send_hdr(self, k: int, op: int, a0: int = 0, a3: int = 0) -> None:
        self._seq[k] += 1
        self._hdr[0], self._hdr[1], self._hdr[2], self._hdr[3] = op, a0, self._seq[k], a3
        dist.send(self._hdr, dst=self._rank(k))
        self.messages += 1

    # ---- data plane ---------------------------------------------------------------------------
    @_locked
    def pull(self) -> Dict[str, torch.Tensor]:
        """reference ``sync_op`` (worker.py:81-85,131): fetch every trainable global variable."""
        out: Dict[str, torch.Tensor] = {}
        for k, lay in enumerate(self.layouts):
            if lay.numel == 0:
                continue
            self._send_hdr(k, OP_PULL)
            buf = torch.empty(lay.numel, dtype=torch.float32)
            dist.recv(buf, src=self._rank(k))
            for v in lay.float_vars:
                o = lay.offsets[v.name]
                out[v.name] = buf[o:o + v.numel].view(*v.shape)
        return out

    @_locked
    def push(self, grads: Dict[str, torch.Tensor], lr: float) -> None:
        """reference ``train_op`` apply half (worker.py:75-79): send gradients, ps applies."""
        for k, lay in enumerate(self.layouts):
            if lay.numel == 0:
                continue
            buf = torch.empty(lay.numel, dtype=torch.float32)
            for v in lay.float_vars:
                o = lay.offsets[v.name]
                buf[o:o + v.numel] = grads[v.name].reshape(-1)
            self._send_hdr(k, OP_PUSH, _float_to_bits(lr))
            dist.send(buf, dst=self._rank(k))

    @_locked
    def push_opt(self, grads: Dict[str, torch.Tensor], lr: float, adam: Optional[dict] = None, sync: bool = False,
                 timeout_s: float = 60.0) -> Optional[int]:
        """Push with the optional modes (module docstring).  Asynchronous: returns None.  Synchronous: blocks
        until every worker's push of this round has been applied as one mean update and returns the global
        step AFTER it; raises ``SyncAborted`` if the round was released without an update or timed out."""
        import datetime
        use_adam = adam is not None
        hyper = [adam.get("beta1", 0.9), adam.get("beta2", 0.999), adam.get("eps", 1e-8)] if use_adam else [0.0, 0.0, 0.0]
        flags = (F_ADAM if use_adam else 0) | (F_SYNC if sync else 0)
        active = []
        for k, lay in enumerate(self.layouts):
            if not lay.trainable and not (sync and lay.has_step):
                continue
            desc = lay.push_descriptor(use_adam)
            nv = len(desc) // 4
            payload = torch.empty(3 + sum(desc[3::4]), dtype=torch.float32)
            payload[0], payload[1], payload[2] = hyper
            c = 3
            for v in lay.trainable:
                payload[c:c + v.numel] = grads[v.name].reshape(-1)
                c += v.numel
            self._send_hdr(k, OP_PUSH_OPT, _float_to_bits(lr), flags | (nv << 8))
            dist.send(torch.tensor(desc if desc else [0], dtype=torch.int64), dst=self._rank(k))
            dist.send(payload, dst=self._rank(k))
            active.append(k)
        if not sync:
            return None
        new_step, aborted = None, False
        for k in active:
            t = torch.zeros(1, dtype=torch.int64)
            try:
                dist.irecv(t, src=self._rank(k)).wait(timeout=datetime.timedelta(seconds=timeout_s))
            except RuntimeError as e:
                raise SyncAborted(f"no answer from ps:{k} within {timeout_s:.0f} s: {str(e).splitlines()[0]}")
            if int(t.item()) == SYNC_ABORTED:
                aborted = True
            elif k == self.step_task:
                new_step = int(t.item())
        if aborted or new_step is None:
            raise SyncAborted("the round was released without an update (a peer left or stop was requested)")
        return new_step

    @_locked
    def read_step(self) -> int:
        self._send_hdr(self.step_task, OP_STEP_READ)
        t = torch.zeros(1, dtype=torch.int64)
        dist.recv(t, src=self._rank(self.step_task))
        return int(t.item())

    @_locked
    def increment_step(self, inc: int = 1) -> int:
        """reference ``counter_op`` (worker.py:32,141); returns the pre-increment value."""
        self._send_hdr(self.step_task, OP_STEP_INC, inc)
        t = torch.zeros(1, dtype=torch.int64)
        dist.recv(t, src=self._rank(self.step_task))
        return int(t.item())

    # ---- control plane --------------------------------------------------------------------------
    def initialize(self, rng):
        vals = {v.name: v.initializer(v.shape, v.dtype, rng) for v in self.variables}
        self.assign_many(vals)

    @_locked
    def assign_many(self, values):
        step = values.get(self.global_step_name)
        for k, lay in enumerate(self.layouts):
            self._send_hdr(k, OP_ALLOC, lay.numel)
            buf = torch.zeros(max(lay.numel, 1) + 1, dtype=torch.float32)
            for v in lay.float_vars:
                o = lay.offsets[v.name]
                buf[o:o + v.numel] = torch.from_numpy(np.ascontiguousarray(values[v.name], dtype=np.float32).reshape(-1))
            s = int(np.asarray(step).reshape(())) if (lay.has_step and step is not None) else -1
            # every task keeps its own Adam step count (each sees every push once): restore it everywhere
            adam_t = values.get(self.adam_t_name)
            self._send_hdr(k, OP_ASSIGN, s, int(np.asarray(adam_t).reshape(())) if adam_t is not None else -1)
            dist.send(buf, dst=self._rank(k))

    @_locked
    def snapshot(self, names=None):
        pulled = self.pull()
        out = {n: t.numpy().copy() for n, t in pulled.items()}
        out[self.global_step_name] = np.array(self.read_step(), dtype=np.int32)
        if any(lay.has_adam_t for lay in self.layouts):
            self._send_hdr(self.step_task, OP_ADAM_T_READ)
            t = torch.zeros(1, dtype=torch.int64)
            dist.recv(t, src=self._rank(self.step_task))
            out[self.adam_t_name] = np.array(int(t.item()), dtype=np.int32)
        return {n: out[n] for n in names} if names is not None else out

    @_locked
    def mark_ready(self):
        for k in range(self.num_ps):
            self._send_hdr(k, OP_SET_READY)

    @_locked
    def report_uninitialized_variables(self):
        for k, lay in enumerate(self.layouts):
            self._send_hdr(k, OP_IS_READY)
            t = torch.zeros(1, dtype=torch.int64)
            dist.recv(t, src=self._rank(k))
            if int(t.item()) == 0:
                return [v.name for v in self.variables]
        return []

    @_locked
    def request_stop(self):
        for k in range(self.num_ps):      # every ps task runs its own round barrier in synchronous mode
            self._send_hdr(k, OP_STOP_REQ)

    @_locked
    def stop_requested(self) -> bool:
        self._send_hdr(0, OP_STOP_QUERY)
        t = torch.zeros(1, dtype=torch.int64)
        dist.recv(t, src=self._rank(0))
        return bool(t.item())

    @_locked
    def shutdown_servers(self, force: bool = False) -> None:
        """Tell every ps this worker is done (the ps exits once all workers said so)."""
        for k in range(self.num_ps):
            self._send_hdr(k, OP_SHUTDOWN, 1 if force else 0)
