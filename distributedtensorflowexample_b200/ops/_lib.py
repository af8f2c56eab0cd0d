"""ctypes loaders for the two in-tree native libraries.

The CUDA library must be present on a GPU box: ``cuda_lib()`` raises (never silently falls
back to eager PyTorch) when it cannot be loaded, so a missing build is a loud failure.
"""
from __future__ import annotations

import ctypes
import os
import threading
from pathlib import Path

from .. import _build

_lock = threading.Lock()
_cuda = None
_host = None


class NativeError(RuntimeError):
    pass


def _load(path: Path) -> ctypes.CDLL:
    if not path.exists():
        raise NativeError(f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                          f"or `python -m distributedtensorflowexample_b200._build`")
    return ctypes.CDLL(str(path), mode=ctypes.RTLD_GLOBAL)


def host_lib(build_if_missing: bool = True) -> ctypes.CDLL:
    """libdtfe_host.so (crc32c / sstable / bundle / events / idx / rendezvous). CPU only."""
    global _host
    with _lock:
        if _host is None:
            if build_if_missing and os.environ.get("DTFE_NO_AUTOBUILD") != "1":
                _build.build_host()
            _host = _load(_build.HOST_LIB)
            _host.dtfe_host_last_error.restype = ctypes.c_char_p
        return _host


def cuda_lib(build_if_missing: bool = True) -> ctypes.CDLL:
    """libdtfe_cuda.so (all sm_100a kernels)."""
    global _cuda
    with _lock:
        if _cuda is None:
            if build_if_missing and not _build.CUDA_LIB.exists() and os.environ.get("DTFE_NO_AUTOBUILD") != "1":
                _build.build_cuda()
            # DTFE_CUDA_LIB: load another build of the library (A/B timing of kernel variants)
            override = os.environ.get("DTFE_CUDA_LIB")
            _cuda = _load(_build.CUDA_LIB if not override else type(_build.CUDA_LIB)(override))
            _cuda.dtfe_last_error.restype = ctypes.c_char_p
            _cuda.dtfe_mlp2_step_workspace_words.restype = ctypes.c_size_t
        return _cuda


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = cuda_lib().dtfe_last_error().decode(errors="replace")
        raise NativeError(f"{what or 'dtfe call'} failed (rc={rc}): {msg}")


def check_host(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = host_lib().dtfe_host_last_error().decode(errors="replace")
        raise NativeError(f"{what or 'dtfe host call'} failed (rc={rc}): {msg}")
