"""ctypes bindings of ``libdtfe_host.so``: crc32c, SSTable, tensor-bundle checkpoints,
tfevents, idx reader.  (formats in Appendix B/C/D.)"""
from __future__ import annotations

import ctypes
import struct
from ctypes import POINTER, c_char_p, c_double, c_float, c_int, c_int64, c_size_t, c_uint32, c_uint64, c_void_p
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

from ..ops._lib import host_lib, check_host, NativeError

_configured = False

# TensorFlow DataType enum values used by the bundle format
DT_FLOAT, DT_DOUBLE, DT_INT32, DT_INT64, DT_BFLOAT16 = 1, 2, 3, 9, 14
_NP_TO_DT = {np.dtype(np.float32): DT_FLOAT, np.dtype(np.float64): DT_DOUBLE, np.dtype(np.int32): DT_INT32,
             np.dtype(np.int64): DT_INT64}
_DT_TO_NP = {v: k for k, v in _NP_TO_DT.items()}
_DT_TO_NP[DT_BFLOAT16] = np.dtype(np.uint16)   # raw bits; callers reinterpret


def _lib():
    global _configured
    L = host_lib()
    if not _configured:
        L.dtfe_crc32c.argtypes = [c_void_p, c_size_t]; L.dtfe_crc32c.restype = c_uint32
        L.dtfe_crc32c_extend.argtypes = [c_uint32, c_void_p, c_size_t]; L.dtfe_crc32c_extend.restype = c_uint32
        L.dtfe_crc32c_masked.argtypes = [c_void_p, c_size_t]; L.dtfe_crc32c_masked.restype = c_uint32
        L.dtfe_bundle_writer_new.argtypes = []; L.dtfe_bundle_writer_new.restype = c_void_p
        L.dtfe_bundle_writer_add.argtypes = [c_void_p, c_char_p, c_int, POINTER(c_int64), c_int, c_void_p, c_size_t]
        L.dtfe_bundle_writer_finish.argtypes = [c_void_p, c_char_p]
        L.dtfe_bundle_writer_abort.argtypes = [c_void_p]; L.dtfe_bundle_writer_abort.restype = None
        L.dtfe_bundle_reader_open.argtypes = [c_char_p]; L.dtfe_bundle_reader_open.restype = c_void_p
        L.dtfe_bundle_reader_count.argtypes = [c_void_p]
        L.dtfe_bundle_reader_entry.argtypes = [c_void_p, c_int, POINTER(c_int64), POINTER(c_int64)]
        L.dtfe_bundle_reader_entry.restype = c_char_p
        L.dtfe_bundle_reader_copy.argtypes = [c_void_p, c_int, c_void_p, c_size_t]
        L.dtfe_bundle_reader_close.argtypes = [c_void_p]; L.dtfe_bundle_reader_close.restype = None
        L.dtfe_sstable_write.argtypes = [c_char_p, POINTER(c_char_p), POINTER(c_size_t), POINTER(c_char_p),
                                         POINTER(c_size_t), c_int, c_size_t]
        L.dtfe_sstable_read.argtypes = [c_char_p, POINTER(c_void_p), POINTER(c_size_t), POINTER(c_int)]
        L.dtfe_host_free.argtypes = [c_void_p]; L.dtfe_host_free.restype = None
        L.dtfe_events_open.argtypes = [c_char_p]; L.dtfe_events_open.restype = c_void_p
        L.dtfe_events_write_version.argtypes = [c_void_p, c_double]
        L.dtfe_events_write_scalars.argtypes = [c_void_p, c_double, c_int64, POINTER(c_char_p), POINTER(c_float), c_int]
        L.dtfe_events_write_scalars_batch.argtypes = [c_void_p, c_double, POINTER(c_int64), POINTER(c_char_p),
                                                      POINTER(c_float), c_int, c_int]
        L.dtfe_events_write_pairs_i32.argtypes = [c_void_p, c_double, c_void_p, c_void_p, c_void_p, c_int, c_char_p, c_char_p]
        L.dtfe_events_write_session_log.argtypes = [c_void_p, c_double, c_int64, c_int, c_char_p, c_char_p]
        L.dtfe_events_write_graph.argtypes = [c_void_p, c_double, c_void_p, c_size_t]
        L.dtfe_events_flush.argtypes = [c_void_p]
        L.dtfe_events_record_count.argtypes = [c_void_p]; L.dtfe_events_record_count.restype = c_uint64
        L.dtfe_events_close.argtypes = [c_void_p]; L.dtfe_events_close.restype = None
        L.dtfe_idx_read.argtypes = [c_char_p, POINTER(c_int), POINTER(c_int64), POINTER(c_void_p), POINTER(c_size_t)]
        L.dtfe_u8_to_unit_float.argtypes = [c_void_p, c_void_p, c_size_t]; L.dtfe_u8_to_unit_float.restype = None
        _configured = True
    return L


# ---- crc32c -------------------------------------------------------------------------------
def crc32c(data: bytes) -> int:
    return int(_lib().dtfe_crc32c(data, len(data)))


def crc32c_extend(crc: int, data: bytes) -> int:
    """crc32c of (already-summed prefix + data): ``crc32c(a + b) == crc32c_extend(crc32c(a), b)``."""
    return int(_lib().dtfe_crc32c_extend(crc, data, len(data)))


def crc32c_masked(data: bytes) -> int:
    return int(_lib().dtfe_crc32c_masked(data, len(data)))


# ---- SSTable ------------------------------------------------------------------------------
def sstable_write(path: str, items: Sequence[Tuple[bytes, bytes]], block_size: int = 256 * 1024) -> None:
    n = len(items)
    keys = (c_char_p * n)(*[k for k, _ in items])
    vals = (c_char_p * n)(*[v for _, v in items])
    klens = (c_size_t * n)(*[len(k) for k, _ in items])
    vlens = (c_size_t * n)(*[len(v) for _, v in items])
    check_host(_lib().dtfe_sstable_write(path.encode(), keys, klens, vals, vlens, n, block_size), "sstable_write")


def sstable_read(path: str) -> List[Tuple[bytes, bytes]]:
    out, n, cnt = c_void_p(), c_size_t(), c_int()
    check_host(_lib().dtfe_sstable_read(path.encode(), ctypes.byref(out), ctypes.byref(n), ctypes.byref(cnt)),
               "sstable_read")
    try:
        flat = ctypes.string_at(out.value, n.value)
    finally:
        _lib().dtfe_host_free(out)
    items, p = [], 0
    for _ in range(cnt.value):
        kl, = struct.unpack_from("<I", flat, p); p += 4
        k = flat[p:p + kl]; p += kl
        vl, = struct.unpack_from("<I", flat, p); p += 4
        v = flat[p:p + vl]; p += vl
        items.append((k, v))
    return items


# ---- tensor bundle ------------------------------------------------------------------------
def bundle_write(prefix: str, tensors: Dict[str, np.ndarray], dtypes: Optional[Dict[str, int]] = None) -> None:
    """Write ``<prefix>.index`` + ``<prefix>.data-00000-of-00001``.  ``dtypes`` overrides the TF dtype
    code per tensor (e.g. DT_BFLOAT16 for uint16 raw bits)."""
    L = _lib()
    w = L.dtfe_bundle_writer_new()
    try:
        for name, arr in tensors.items():
            a = np.asarray(arr)
            if not a.flags.c_contiguous:
                a = a.copy(order="C")
            dt = (dtypes or {}).get(name)
            if dt is None:
                if a.dtype not in _NP_TO_DT:
                    raise TypeError(f"{name}: dtype {a.dtype} not supported by the bundle writer")
                dt = _NP_TO_DT[a.dtype]
            shape = (c_int64 * max(a.ndim, 1))(*a.shape)
            check_host(L.dtfe_bundle_writer_add(w, name.encode(), dt, shape, a.ndim, a.ctypes.data, a.nbytes),
                       f"bundle add {name}")
    except Exception:
        L.dtfe_bundle_writer_abort(w)
        raise
    check_host(L.dtfe_bundle_writer_finish(w, prefix.encode()), "bundle finish")


def bundle_read(prefix: str) -> Dict[str, np.ndarray]:
    L = _lib()
    r = L.dtfe_bundle_reader_open(prefix.encode())
    if not r:
        raise NativeError(f"cannot read checkpoint {prefix}: " + L.dtfe_host_last_error().decode())
    out: Dict[str, np.ndarray] = {}
    try:
        for i in range(L.dtfe_bundle_reader_count(r)):
            meta = (c_int64 * 5)()
            shape = (c_int64 * 8)()
            name = L.dtfe_bundle_reader_entry(r, i, meta, shape).decode()
            dt, rank, nbytes = int(meta[0]), int(meta[1]), int(meta[2])
            if dt not in _DT_TO_NP:
                raise NativeError(f"{name}: TF dtype code {dt} not supported")
            arr = np.empty([int(shape[d]) for d in range(rank)], dtype=_DT_TO_NP[dt])
            if arr.nbytes != nbytes:
                raise NativeError(f"{name}: entry says {nbytes} bytes but shape/dtype give {arr.nbytes}")
            check_host(L.dtfe_bundle_reader_copy(r, i, arr.ctypes.data, arr.nbytes), f"bundle copy {name}")
            out[name] = arr
    finally:
        L.dtfe_bundle_reader_close(r)
    return out


def bundle_entries(prefix: str) -> List[dict]:
    """Index metadata only: name, dtype code, shape, offset, size, masked crc (layout goldens)."""
    L = _lib()
    r = L.dtfe_bundle_reader_open(prefix.encode())
    if not r:
        raise NativeError(f"cannot read checkpoint {prefix}: " + L.dtfe_host_last_error().decode())
    try:
        out = []
        for i in range(L.dtfe_bundle_reader_count(r)):
            meta = (c_int64 * 5)(); shape = (c_int64 * 8)()
            name = L.dtfe_bundle_reader_entry(r, i, meta, shape).decode()
            out.append(dict(name=name, dtype=int(meta[0]), shape=[int(shape[d]) for d in range(int(meta[1]))],
                            size=int(meta[2]), offset=int(meta[3]), crc32c=int(meta[4]) & 0xffffffff))
        return out
    finally:
        L.dtfe_bundle_reader_close(r)


# ---- tfevents -----------------------------------------------------------------------------
SESSION_START, SESSION_STOP, SESSION_CHECKPOINT = 1, 2, 3


class EventsFile:
    def __init__(self, path: str):
        self._L = _lib()
        self._h = self._L.dtfe_events_open(path.encode())
        if not self._h:
            raise NativeError(self._L.dtfe_host_last_error().decode())
        self.path = path

    def write_version(self, wall_time: float) -> None:
        check_host(self._L.dtfe_events_write_version(self._h, wall_time), "events version")

    def write_scalars(self, wall_time: float, step: int, tags: Sequence[str], values: Sequence[float]) -> None:
        n = len(tags)
        T = (c_char_p * n)(*[t.encode() for t in tags])
        V = (c_float * n)(*[float(v) for v in values])
        check_host(self._L.dtfe_events_write_scalars(self._h, wall_time, int(step), T, V, n), "events scalars")

    def write_scalars_batch(self, wall_time: float, steps: Sequence[int], tags: Sequence[str], rows) -> None:
        """One event per step: ``rows[i]`` holds the ``len(tags)`` scalar values of ``steps[i]``."""
        n, nt = len(steps), len(tags)
        if n == 0:
            return
        T = (c_char_p * nt)(*[t.encode() for t in tags])
        S = (c_int64 * n)(*[int(x) for x in steps])
        V = (c_float * (n * nt))(*[float(v) for row in rows for v in row])
        check_host(self._L.dtfe_events_write_scalars_batch(self._h, wall_time, S, T, V, n, nt), "events scalars batch")

    def write_pairs_i32(self, wall_time: float, steps: np.ndarray, v0: np.ndarray, v1: np.ndarray, tag0: str, tag1: str) -> None:
        """One event per step with two scalars, from contiguous int32 / float32 / float32 arrays (no per-value
        marshalling; the call releases the GIL)."""
        n = int(steps.shape[0])
        if n == 0:
            return
        assert steps.dtype == np.int32 and v0.dtype == np.float32 and v1.dtype == np.float32
        assert steps.flags.c_contiguous and v0.flags.c_contiguous and v1.flags.c_contiguous and len(v0) == n == len(v1)
        check_host(self._L.dtfe_events_write_pairs_i32(self._h, wall_time, steps.ctypes.data, v0.ctypes.data,
                                                       v1.ctypes.data, n, tag0.encode(), tag1.encode()), "events pairs")

    def write_session_log(self, wall_time: float, step: int, status: int, checkpoint_path: str = "", msg: str = "") -> None:
        check_host(self._L.dtfe_events_write_session_log(self._h, wall_time, int(step), status,
                                                         checkpoint_path.encode(), msg.encode()), "events session_log")

    def write_graph(self, wall_time: float, graph_def: bytes) -> None:
        check_host(self._L.dtfe_events_write_graph(self._h, wall_time, graph_def, len(graph_def)), "events graph")

    def flush(self) -> None:
        if self._h:
            check_host(self._L.dtfe_events_flush(self._h), "events flush")

    @property
    def record_count(self) -> int:
        return int(self._L.dtfe_events_record_count(self._h))

    def close(self) -> None:
        if self._h:
            self._L.dtfe_events_close(self._h)
            self._h = None


# ---- idx ----------------------------------------------------------------------------------
def read_idx(path: str) -> np.ndarray:
    L = _lib()
    nd, dims, data, nbytes = c_int(), (c_int64 * 4)(), c_void_p(), c_size_t()
    check_host(L.dtfe_idx_read(path.encode(), ctypes.byref(nd), dims, ctypes.byref(data), ctypes.byref(nbytes)),
               "idx read")
    try:
        buf = np.frombuffer(ctypes.string_at(data.value, nbytes.value), dtype=np.uint8)
    finally:
        L.dtfe_host_free(data)
    return buf.reshape([int(dims[i]) for i in range(nd.value)])
