"""Supervisor: chief election, init-or-restore, readiness wait, periodic services, stop
coordination (reference worker.py:107-123; TF1 tf.train.Supervisor / SessionManager /
Coordinator - SURVEY C16/T12, 3.3, 3.5, 5.3, 5.4).

The "session" a Supervisor manages here is a parameter client (``parallel.client.ParamClient``):
an object that can initialise, assign, read and snapshot the global variables wherever they live
(peer HBM shards or the gloo ps process).

Lifecycle, same as TF1:
  chief      : restore the newest checkpoint in ``logdir`` if there is one, else run ``init_op``;
               run ``local_init_op``; check ``ready_op`` (no uninitialised global variable);
               write ``graph.pbtxt``; add the graph + SessionLog START to the summary writer; start
               the checkpoint thread (first save fires immediately -> ``model.ckpt-<step>``) and the
               ``global_step/sec`` thread.  No summary thread (the reference passes summary_op=None).
  non-chief  : loop { local_init_op; ready_op; if not ready: sleep recovery_wait_secs }.
  exit       : the context manager requests stop on exception, joins the threads; no final save.
"""
from __future__ import annotations

import contextlib
import os
import threading
import time
import traceback
from typing import Callable, List, Optional

import numpy as np

from ..utils import checkpoint as ckpt
from ..utils.events import FileWriter, Summary, TAG_STEPS_PER_SEC
from ..utils.native_io import SESSION_CHECKPOINT, SESSION_START
from ..utils.tracing import get_tracer


class Coordinator:
    def __init__(self) -> None:
        self._stop = threading.Event()
        self._exc: Optional[BaseException] = None
        self._lock = threading.Lock()

    def request_stop(self, ex: Optional[BaseException] = None) -> None:
        with self._lock:
            if ex is not None and self._exc is None:
                self._exc = ex
        self._stop.set()

    def should_stop(self) -> bool:
        return self._stop.is_set()

    def wait_for_stop(self, timeout: Optional[float] = None) -> bool:
        return self._stop.wait(timeout)

    def raise_requested_exception(self) -> None:
        if self._exc is not None:
            raise self._exc


class _LoopThread(threading.Thread):
    """TF1 ``LooperThread``: ``start_loop`` once, then ``run_loop`` every ``period`` seconds."""

    def __init__(self, coord: Coordinator, period: float, name: str):
        super().__init__(name=name, daemon=True)
        self._coord, self._period = coord, period

    def start_loop(self) -> None: ...
    def run_loop(self) -> None: ...

    def run(self) -> None:
        try:
            self.start_loop()
            while not self._coord.should_stop():
                t0 = time.time()
                self.run_loop()
                if self._coord.wait_for_stop(max(0.0, self._period - (time.time() - t0))):
                    break
        except BaseException as e:  # a failing service stops training, like TF1's Coordinator
            traceback.print_exc()
            self._coord.request_stop(e)


class Supervisor:
    def __init__(self, is_chief: bool, logdir: str, saver: Optional[ckpt.Saver], client,
                 summary_writer: Optional[FileWriter] = None, save_model_secs: float = 30,
                 save_summaries_secs: float = 30, init_op: Optional[Callable[[], None]] = None,
                 local_init_op: Optional[Callable[[], None]] = None,
                 ready_op: Optional[Callable[[], List[str]]] = None, recovery_wait_secs: float = 30,
                 checkpoint_basename: str = "model.ckpt", graph_text: str = "", global_step_name: str = "global/global_step"):
        self.is_chief = is_chief
        self.logdir = logdir
        self.saver = saver
        self.client = client
        self.summary_writer = summary_writer
        self.save_model_secs = save_model_secs
        self.save_summaries_secs = save_summaries_secs
        self.init_op = init_op
        self.local_init_op = local_init_op
        self.ready_op = ready_op or (lambda: client.report_uninitialized_variables())
        self.recovery_wait_secs = recovery_wait_secs
        self.save_path = os.path.join(logdir, checkpoint_basename)
        self.graph_text = graph_text
        self.global_step_name = global_step_name
        self.coord = Coordinator()
        self._threads: List[threading.Thread] = []
        self.restored_from: Optional[str] = None

    # ---- stop plumbing --------------------------------------------------------------------
    def should_stop(self) -> bool:
        return self.coord.should_stop()

    def request_stop(self, ex: Optional[BaseException] = None) -> None:
        self.coord.request_stop(ex)

    # ---- session bring-up -------------------------------------------------------------------
    def _prepare_chief(self) -> None:
        latest = ckpt.latest_checkpoint(self.logdir) if self.saver is not None else None
        if latest is not None:
            values = self.saver.restore(latest)
            self.client.assign_many(values)
            self.restored_from = latest
        elif self.init_op is not None:
            self.init_op()
        if self.local_init_op is not None:
            self.local_init_op()
        self.client.mark_ready()
        not_ready = self.ready_op()
        if not_ready:
            raise RuntimeError(f"Init operations did not make model ready; uninitialized: {not_ready}")
        os.makedirs(self.logdir, exist_ok=True)
        with open(os.path.join(self.logdir, "graph.pbtxt"), "w") as f:
            f.write(self.graph_text)
        if self.summary_writer is not None:
            self.summary_writer.add_graph(self.graph_text)
            self.summary_writer.add_session_log(SESSION_START, 0)

    def _wait_non_chief(self) -> None:
        while not self.coord.should_stop():
            if self.local_init_op is not None:
                self.local_init_op()
            try:
                if not self.ready_op():
                    return
            except Exception:
                pass
            if self.coord.wait_for_stop(self.recovery_wait_secs):
                break

    def _start_services(self) -> None:
        if self.saver is not None and self.save_model_secs:
            self._threads.append(_CheckpointThread(self))
        if self.summary_writer is not None and self.save_summaries_secs:
            self._threads.append(_StepCounterThread(self))
        for t in self._threads:
            t.start()

    @contextlib.contextmanager
    def managed_session(self):
        """Yields the parameter client once the model is ready (TF1 ``sv.managed_session``)."""
        try:
            if self.is_chief:
                self._prepare_chief()
                self._start_services()
            else:
                self._wait_non_chief()
            yield self.client
        except BaseException as e:
            # SystemExit from the signal handlers lands here too (reference main.py:51-52)
            self.coord.request_stop(e if not isinstance(e, (SystemExit, KeyboardInterrupt)) else None)
            raise
        finally:
            self.coord.request_stop()
            for t in self._threads:
                t.join(timeout=10)
            if self.summary_writer is not None:
                self.summary_writer.flush()

    # ---- services ---------------------------------------------------------------------------
    def save_checkpoint(self) -> str:
        tr = get_tracer()
        with tr.span("checkpoint.snapshot", "checkpoint"):
            values = self.client.snapshot(self.saver.var_list)
        step = int(np.asarray(values[self.global_step_name]).reshape(()))
        with tr.span("checkpoint.write", "checkpoint", step=step):
            path = self.saver.save(values, self.save_path, global_step=step)
        if self.summary_writer is not None:
            self.summary_writer.add_session_log(SESSION_CHECKPOINT, step, checkpoint_path=path)
        return path


class _CheckpointThread(_LoopThread):
    """TF1 SVTimerCheckpointThread: every save_model_secs, first fire immediately."""

    def __init__(self, sv: Supervisor):
        super().__init__(sv.coord, sv.save_model_secs, "sv-checkpoint")
        self._sv = sv

    def run_loop(self) -> None:
        self._sv.save_checkpoint()


class _StepCounterThread(_LoopThread):
    """TF1 SVStepCounterThread: global steps per second, tag ``global/global_step/sec``."""

    def __init__(self, sv: Supervisor):
        super().__init__(sv.coord, sv.save_summaries_secs, "sv-stepcounter")
        self._sv = sv
        self._last_time = 0.0
        self._last_step = 0

    def start_loop(self) -> None:
        self._last_time = time.time()
        self._last_step = self._sv.client.read_step()

    def run_loop(self) -> None:
        step = self._sv.client.read_step()
        now = time.time()
        dt = now - self._last_time
        if dt <= 0:
            return
        rate = (step - self._last_step) / dt
        self._last_time, self._last_step = now, step
        if self._sv.summary_writer is not None:
            self._sv.summary_writer.add_summary(Summary([(TAG_STEPS_PER_SEC, rate)]), step)
