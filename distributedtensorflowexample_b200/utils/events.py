"""Summaries and the TensorBoard event writer (reference worker.py:87-96,139; TF1
tf.summary.scalar / merge_all / FileWriter - SURVEY C13/T14, Appendix C).

``FileWriter(logdir)`` creates ``events.out.tfevents.<unix_secs>.<hostname>`` immediately (like TF1,
the file exists as soon as the writer is constructed) and appends records through the native
writer.  Writes are buffered and flushed every ``flush_secs`` (TF1 default 120 s) by a daemon
thread, so the per-step ``add_summary`` the reference does (worker.py:139) is off the critical path.
"""
from __future__ import annotations

import os
import socket
import threading
import time
from typing import Optional, Sequence, Tuple

from . import native_io

# Tags as TF1 would name them for the reference graph: the scalars are created inside
# variable_scope('local'); the name scope 'accuracy' is already taken (worker.py:88) so the second
# use is uniquified to 'accuracy_1' (SURVEY T14).
TAG_LOSS = "local/loss"
TAG_ACCURACY = "local/accuracy_1"
TAG_STEPS_PER_SEC = "global/global_step/sec"
TAG_TEST_ACCURACY = "local/test_accuracy"     # additive: the reference prints the evaluation, it does not log it


class Summary:
    """A merged scalar summary: ordered (tag, value) pairs (``tf.summary.merge_all`` output)."""

    def __init__(self, values: Optional[Sequence[Tuple[str, float]]] = None):
        self.values = list(values or [])

    def add(self, tag: str, value: float) -> "Summary":
        self.values.append((tag, float(value)))
        return self


def training_summary(loss: float, accuracy: float) -> Summary:
    return Summary([(TAG_LOSS, loss), (TAG_ACCURACY, accuracy)])


class FileWriter:
    def __init__(self, logdir: str, flush_secs: float = 120.0, filename_suffix: str = ""):
        os.makedirs(logdir, exist_ok=True)
        self.logdir = logdir
        now = time.time()
        name = f"events.out.tfevents.{int(now):010d}.{socket.gethostname()}{filename_suffix}"
        self.path = os.path.join(logdir, name)
        self._f = native_io.EventsFile(self.path)
        self._f.write_version(now)
        self._f.flush()
        self._closed = False
        self._flush_secs = flush_secs
        self._wake = threading.Event()
        self._thread = threading.Thread(target=self._flusher, name="events-flush", daemon=True)
        self._thread.start()

    def _flusher(self) -> None:
        while not self._wake.wait(self._flush_secs):
            try:
                self._f.flush()
            except Exception:
                return

    def add_summary(self, summary: Summary, global_step: int, wall_time: Optional[float] = None) -> None:
        tags = [t for t, _ in summary.values]
        vals = [v for _, v in summary.values]
        self._f.write_scalars(time.time() if wall_time is None else wall_time, int(global_step), tags, vals)

    def add_training_summaries(self, global_steps, losses, accuracies, wall_time: Optional[float] = None) -> None:
        """The per-step loss/accuracy summaries (reference worker.py:92-95,139) of a whole multi-step launch in
        one native call; byte-identical to calling ``add_summary(training_summary(l, a), s)`` per step."""
        self._f.write_scalars_batch(time.time() if wall_time is None else wall_time, list(global_steps),
                                    [TAG_LOSS, TAG_ACCURACY], list(zip(losses, accuracies)))

    def add_training_arrays(self, global_steps, losses, accuracies, wall_time: Optional[float] = None) -> None:
        """``add_training_summaries`` for numpy arrays (int32 steps, float32 loss/accuracy), written by the native
        library straight from the arrays."""
        self._f.write_pairs_i32(time.time() if wall_time is None else wall_time, global_steps, losses, accuracies,
                                TAG_LOSS, TAG_ACCURACY)

    def add_session_log(self, status: int, global_step: int = 0, checkpoint_path: str = "", msg: str = "") -> None:
        self._f.write_session_log(time.time(), int(global_step), status, checkpoint_path, msg)

    def add_graph(self, graph_text: str) -> None:
        self._f.write_graph(time.time(), graph_text.encode())

    def flush(self) -> None:
        if not self._closed:
            self._f.flush()

    def close(self) -> None:
        if self._closed:
            return
        self._closed = True
        self._wake.set()
        self._f.flush()
        self._f.close()
