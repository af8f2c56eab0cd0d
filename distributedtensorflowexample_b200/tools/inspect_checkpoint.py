"""``python -m distributedtensorflowexample_b200.tools.inspect_checkpoint <prefix | logdir>`` - list the
tensors of a TF1 V2 checkpoint (what TF's ``inspect_checkpoint`` tool prints), optionally with values."""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np

from ..utils import checkpoint as ckpt
from ..utils import native_io


def resolve_prefix(path: str) -> str:
    if os.path.isdir(path):
        latest = ckpt.latest_checkpoint(path)
        if latest is None:
            raise FileNotFoundError(f"no checkpoint state file in {path}")
        return latest
    for suffix in (".index", ".data-00000-of-00001"):
        if path.endswith(suffix):
            return path[:-len(suffix)]
    return path


def describe(prefix: str, tensor: str = "", all_tensors: bool = False) -> str:
    tensors = native_io.bundle_read(prefix)
    lines = [f"checkpoint: {prefix}"]
    if tensor:
        if tensor not in tensors:
            raise KeyError(f"{tensor} not in checkpoint (has: {', '.join(sorted(tensors))})")
        a = tensors[tensor]
        lines += [f"tensor_name: {tensor}  dtype={a.dtype}  shape={list(a.shape)}", np.array2string(a, threshold=1 << 30)]
        return "\n".join(lines)
    total = 0
    for name in sorted(tensors):
        a = tensors[name]
        total += a.size
        lines.append(f"{name} (DT_{str(a.dtype).upper()}) {list(a.shape)}")
        if all_tensors:
            lines.append(np.array2string(a, threshold=20, edgeitems=3))
    lines.append(f"# {len(tensors)} tensors, {total} elements")
    return "\n".join(lines)


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__)
    ap.add_argument("path", help="checkpoint prefix (model.ckpt-123), one of its files, or a log directory")
    ap.add_argument("--tensor_name", default="", help="print this tensor in full")
    ap.add_argument("--all_tensors", action="store_true", help="print (abbreviated) values of every tensor")
    a = ap.parse_args(argv)
    try:
        print(describe(resolve_prefix(a.path), a.tensor_name, a.all_tensors))
    except (FileNotFoundError, KeyError, native_io.NativeError) as e:
        print(f"inspect_checkpoint: {e}", file=sys.stderr)
        return 1
    return 0


if __name__ == "__main__":
    sys.exit(main())
