"""Tracing: host spans + device (CUDA-event) spans in Chrome trace format, mirrored as NVTX ranges.

The reference has no tracing at all - it even turns device-placement logging off (worker.py:121) and
infers throughput from the wall-clock ``speed`` print (worker.py:143-148); SURVEY 5.1 asks the new
framework for CUDA-event timing around the fused paths and NVTX ranges.  This module is that:

* ``Tracer.span(name)``            host wall-clock span (context manager), also an NVTX range when a CUDA
                                   device is present, so ``ncu --nvtx`` / Nsight can filter on it;
* ``Tracer.device_span(name)``     a pair of CUDA events on the current stream; resolved to microseconds
                                   against the tracer's base event when the trace is written, and shown on a
                                   separate "GPU stream" track next to the host track;
* ``Tracer.instant`` / ``counter`` point events and counter series (e.g. steps/sec);
* ``Tracer.save(path)``            ``{"traceEvents": [...]}`` JSON that chrome://tracing and Perfetto load.

Disabled tracers (the default) hand out a shared no-op context manager: the worker loop pays one
attribute test per span.  ``--trace_file`` turns it on for a task (the task index is appended).
"""
from __future__ import annotations

import contextlib
import json
import os
import threading
import time
from typing import Any, Dict, List, Optional

_NULL = contextlib.nullcontext()


def _nvtx():
    try:
        import torch
        if torch.cuda.is_available():
            return torch.cuda.nvtx
    except Exception:
        pass
    return None


class _Span:
    __slots__ = ("tr", "name", "cat", "args", "t0", "nvtx")

    def __init__(self, tr: "Tracer", name: str, cat: str, args: Optional[Dict[str, Any]]):
        self.tr, self.name, self.cat, self.args = tr, name, cat, args

    def __enter__(self):
        self.nvtx = self.tr._nvtx
        if self.nvtx is not None:
            self.nvtx.range_push(self.name)
        self.t0 = time.perf_counter_ns()
        return self

    def __exit__(self, *exc):
        t1 = time.perf_counter_ns()
        if self.nvtx is not None:
            self.nvtx.range_pop()
        self.tr._add_complete(self.name, self.cat, self.t0, t1 - self.t0, self.args)
        return False


class _DeviceSpan:
    __slots__ = ("tr", "name", "args", "e0", "e1", "stream")

    def __init__(self, tr: "Tracer", name: str, args, stream):
        self.tr, self.name, self.args, self.stream = tr, name, args, stream

    def __enter__(self):
        import torch
        self.e0 = torch.cuda.Event(enable_timing=True)
        self.e1 = torch.cuda.Event(enable_timing=True)
        self.e0.record(self.stream)
        return self

    def __exit__(self, *exc):
        self.e1.record(self.stream)
        self.tr._pending_device.append((self.name, self.e0, self.e1, self.args))
        return False


class Tracer:
    """Collects events in memory (bounded) and writes one Chrome-trace JSON file."""

    def __init__(self, enabled: bool = False, process_name: str = "", pid: Optional[int] = None,
                 max_events: int = 1_000_000, nvtx: bool = True):
        self.enabled = enabled
        self.pid = os.getpid() if pid is None else pid
        self.process_name = process_name
        self.max_events = max_events
        self.dropped = 0
        self._events: List[Dict[str, Any]] = []
        self._pending_device: list = []
        self._lock = threading.Lock()
        self._nvtx = _nvtx() if (enabled and nvtx) else None
        self._t_base_ns = time.perf_counter_ns()
        self._wall_base = time.time()
        self._base_event = None          # CUDA event recorded at the same instant as _t_base_ns (best effort)

    # ---- recording ---------------------------------------------------------------------------
    def span(self, name: str, cat: str = "host", **args):
        if not self.enabled:
            return _NULL
        return _Span(self, name, cat, args or None)

    def device_span(self, name: str, stream=None, **args):
        """Time a region of the current (or given) CUDA stream with events.  No-op without CUDA."""
        if not self.enabled:
            return _NULL
        try:
            import torch
            if not torch.cuda.is_available():
                return _NULL
            if self._base_event is None:
                self._base_event = torch.cuda.Event(enable_timing=True)
                self._base_event.record(stream)
                self._base_event.synchronize()
                self._base_host_ns = time.perf_counter_ns()
        except Exception:
            return _NULL
        return _DeviceSpan(self, name, args or None, stream)

    def instant(self, name: str, cat: str = "host", **args) -> None:
        if self.enabled:
            self._push({"name": name, "cat": cat, "ph": "i", "s": "p", "ts": self._ts(time.perf_counter_ns()),
                        "pid": self.pid, "tid": threading.get_ident() & 0xffff, "args": args or {}})

    def counter(self, name: str, **series) -> None:
        if self.enabled:
            self._push({"name": name, "ph": "C", "ts": self._ts(time.perf_counter_ns()), "pid": self.pid,
                        "args": {k: float(v) for k, v in series.items()}})

    def _ts(self, t_ns: int) -> float:
        return (t_ns - self._t_base_ns) / 1e3

    def _push(self, ev: Dict[str, Any]) -> None:
        with self._lock:
            if len(self._events) >= self.max_events:
                self.dropped += 1
                return
            self._events.append(ev)

    def _add_complete(self, name, cat, t0_ns, dur_ns, args) -> None:
        self._push({"name": name, "cat": cat, "ph": "X", "ts": self._ts(t0_ns), "dur": dur_ns / 1e3,
                    "pid": self.pid, "tid": threading.get_ident() & 0xffff, "args": args or {}})

    # ---- output ------------------------------------------------------------------------------
    def _resolve_device(self) -> None:
        if not self._pending_device:
            return
        pend, self._pending_device = self._pending_device, []
        for name, e0, e1, args in pend:
            try:
                e1.synchronize()
                start_us = self._base_event.elapsed_time(e0) * 1e3 + self._ts(self._base_host_ns)
                dur_us = e0.elapsed_time(e1) * 1e3
            except Exception:
                continue
            self._push({"name": name, "cat": "device", "ph": "X", "ts": start_us, "dur": dur_us, "pid": self.pid,
                        "tid": 1 << 20, "args": args or {}})

    def events(self) -> List[Dict[str, Any]]:
        self._resolve_device()
        with self._lock:
            return list(self._events)

    def summary(self) -> Dict[str, Dict[str, float]]:
        """name -> {count, total_us, mean_us, max_us} over the complete ('X') events."""
        out: Dict[str, Dict[str, float]] = {}
        for ev in self.events():
            if ev.get("ph") != "X":
                continue
            s = out.setdefault(ev["name"], {"count": 0, "total_us": 0.0, "max_us": 0.0})
            s["count"] += 1
            s["total_us"] += ev["dur"]
            s["max_us"] = max(s["max_us"], ev["dur"])
        for s in out.values():
            s["mean_us"] = s["total_us"] / max(1, s["count"])
        return out

    def save(self, path: str) -> Optional[str]:
        if not self.enabled:
            return None
        evs = self.events()
        meta = [{"name": "process_name", "ph": "M", "pid": self.pid, "args": {"name": self.process_name or f"pid {self.pid}"}},
                {"name": "thread_name", "ph": "M", "pid": self.pid, "tid": 1 << 20, "args": {"name": "GPU stream (CUDA events)"}}]
        doc = {"traceEvents": meta + evs, "displayTimeUnit": "ms",
               "otherData": {"wall_clock_base": self._wall_base, "dropped_events": self.dropped}}
        os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
        tmp = f"{path}.tmp{os.getpid()}"
        with open(tmp, "w") as f:
            json.dump(doc, f)
        os.replace(tmp, path)
        return path


_tracer = Tracer(False)


def get_tracer() -> Tracer:
    return _tracer


def enable(process_name: str = "", nvtx: bool = True) -> Tracer:
    """Install an enabled process-wide tracer (idempotent) and return it."""
    global _tracer
    if not _tracer.enabled:
        _tracer = Tracer(True, process_name=process_name, nvtx=nvtx)
    return _tracer


def disable() -> None:
    global _tracer
    _tracer = Tracer(False)


def trace_path_for_task(base: str, job: str, task_index: int) -> str:
    """``run.json`` -> ``run.worker0.json`` so every task of a cluster writes its own file."""
    root, ext = os.path.splitext(base)
    return f"{root}.{job}{task_index}{ext or '.json'}"
