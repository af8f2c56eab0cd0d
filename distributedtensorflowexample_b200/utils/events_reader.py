"""Reader for TensorBoard event files (the counterpart of ``csrc/host/events.cc``).

TFRecord framing - ``u64 length | u32 masked_crc32c(length) | payload | u32 masked_crc32c(payload)`` - with
both checksums verified through the native crc32c, and a decoder for the subset of the ``Event`` proto this
framework (and the reference, worker.py:92-96,111,116) produces: wall_time, step, file_version, scalar
summaries (tag + simple_value), session_log (status, checkpoint_path, msg) and graph_def (length only).
Unknown fields are skipped, so files written by real TensorFlow load too.
"""
from __future__ import annotations

import os
import struct
from dataclasses import dataclass, field
from typing import Dict, Iterator, List, Optional, Tuple

from . import native_io


class EventFileError(RuntimeError):
    pass


@dataclass
class Event:
    wall_time: float = 0.0
    step: int = 0
    file_version: Optional[str] = None
    scalars: List[Tuple[str, float]] = field(default_factory=list)      # (tag, simple_value)
    session_log: Optional[Dict[str, object]] = None                    # {"status": int, "checkpoint_path": str, "msg": str}
    graph_def_bytes: int = 0


def _varint(buf: bytes, p: int) -> Tuple[int, int]:
    v, shift = 0, 0
    while True:
        if p >= len(buf):
            raise EventFileError("truncated varint")
        b = buf[p]
        p += 1
        v |= (b & 0x7F) << shift
        if not b & 0x80:
            return v, p
        shift += 7
        if shift > 63:
            raise EventFileError("varint too long")


def _fields(buf: bytes) -> Iterator[Tuple[int, int, object]]:
    """Yield (field number, wire type, value) of one message; value is int, bytes (len-delimited) or raw bytes."""
    p = 0
    while p < len(buf):
        key, p = _varint(buf, p)
        fno, wt = key >> 3, key & 7
        if wt == 0:
            v, p = _varint(buf, p)
        elif wt == 1:
            v, p = buf[p:p + 8], p + 8
        elif wt == 2:
            n, p = _varint(buf, p)
            v, p = buf[p:p + n], p + n
        elif wt == 5:
            v, p = buf[p:p + 4], p + 4
        else:
            raise EventFileError(f"unsupported wire type {wt}")
        if p > len(buf):
            raise EventFileError("truncated field")
        yield fno, wt, v


def decode_event(payload: bytes) -> Event:
    ev = Event()
    for fno, wt, v in _fields(payload):
        if fno == 1 and wt == 1:
            ev.wall_time = struct.unpack("<d", v)[0]
        elif fno == 2 and wt == 0:
            ev.step = v - (1 << 64) if v >> 63 else v
        elif fno == 3 and wt == 2:
            ev.file_version = v.decode(errors="replace")
        elif fno == 4 and wt == 2:
            ev.graph_def_bytes = len(v)
        elif fno == 5 and wt == 2:                      # Summary
            for f2, w2, val in _fields(v):
                if f2 == 1 and w2 == 2:                 # Summary.Value
                    tag, simple = "", None
                    for f3, w3, x in _fields(val):
                        if f3 == 1 and w3 == 2:
                            tag = x.decode(errors="replace")
                        elif f3 == 2 and w3 == 5:
                            simple = struct.unpack("<f", x)[0]
                    if simple is not None:
                        ev.scalars.append((tag, simple))
        elif fno == 7 and wt == 2:                      # SessionLog
            sl: Dict[str, object] = {"status": 0, "checkpoint_path": "", "msg": ""}
            for f2, w2, x in _fields(v):
                if f2 == 1 and w2 == 0:
                    sl["status"] = x
                elif f2 == 2 and w2 == 2:
                    sl["checkpoint_path"] = x.decode(errors="replace")
                elif f2 == 3 and w2 == 2:
                    sl["msg"] = x.decode(errors="replace")
            ev.session_log = sl
    return ev


def read_records(path: str, verify: bool = True) -> Iterator[bytes]:
    """Payloads of a TFRecord file.  A truncated last record (a writer that is still running) ends the
    iteration silently; a checksum mismatch raises."""
    with open(path, "rb") as f:
        while True:
            head = f.read(12)
            if len(head) < 12:
                return
            (n,), (lcrc,) = struct.unpack("<Q", head[:8]), struct.unpack("<I", head[8:])
            if verify and native_io.crc32c_masked(head[:8]) != lcrc:
                raise EventFileError(f"{path}: length checksum mismatch at offset {f.tell() - 12}")
            body = f.read(n + 4)
            if len(body) < n + 4:
                return
            payload, (pcrc,) = body[:n], struct.unpack("<I", body[n:])
            if verify and native_io.crc32c_masked(payload) != pcrc:
                raise EventFileError(f"{path}: payload checksum mismatch at offset {f.tell() - n - 4}")
            yield payload


def read_events(path: str, verify: bool = True) -> Iterator[Event]:
    """Events of one file, or of every ``events.out.tfevents.*`` file of a directory in name order."""
    if os.path.isdir(path):
        for name in sorted(os.listdir(path)):
            if "tfevents" in name:
                yield from read_events(os.path.join(path, name), verify)
        return
    for payload in read_records(path, verify):
        yield decode_event(payload)


def scalars(path: str, tag: Optional[str] = None) -> List[Tuple[int, float, str, float]]:
    """[(step, wall_time, tag, value)] of every scalar summary (optionally one tag)."""
    return [(e.step, e.wall_time, t, v) for e in read_events(path) for t, v in e.scalars if tag is None or t == tag]
