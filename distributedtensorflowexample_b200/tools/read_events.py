"""``python -m distributedtensorflowexample_b200.tools.read_events <events file | logdir_N>`` - print the
records of TensorBoard event files: scalar summaries, session-log records (start / checkpoint), graph markers."""
from __future__ import annotations

import argparse
import sys

from ..utils import events_reader

_STATUS = {0: "STATUS_UNSPECIFIED", 1: "START", 2: "STOP", 3: "CHECKPOINT"}


def format_events(path: str, tag: str = "", limit: int = 0) -> str:
    out, n = [], 0
    for e in events_reader.read_events(path):
        if e.file_version is not None:
            line = f"{e.wall_time:.3f} file_version {e.file_version}"
        elif e.session_log is not None:
            sl = e.session_log
            line = f"{e.wall_time:.3f} step {e.step} session_log {_STATUS.get(sl['status'], sl['status'])}"
            if sl["checkpoint_path"]:
                line += f" {sl['checkpoint_path']}"
        elif e.graph_def_bytes:
            line = f"{e.wall_time:.3f} graph_def {e.graph_def_bytes} bytes"
        elif e.scalars:
            vals = [(t, v) for t, v in e.scalars if not tag or t == tag]
            if not vals:
                continue
            line = f"{e.wall_time:.3f} step {e.step} " + " ".join(f"{t}={v:.6g}" for t, v in vals)
        else:
            continue
        if tag and not e.scalars:
            continue
        out.append(line)
        n += 1
        if limit and n >= limit:
            break
    return "\n".join(out)


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__)
    ap.add_argument("path")
    ap.add_argument("--tag", default="", help="only scalar summaries with this tag (e.g. local/loss)")
    ap.add_argument("--limit", type=int, default=0, help="stop after this many lines")
    a = ap.parse_args(argv)
    try:
        print(format_events(a.path, a.tag, a.limit))
    except (OSError, events_reader.EventFileError) as e:
        print(f"read_events: {e}", file=sys.stderr)
        return 1
    return 0


if __name__ == "__main__":
    sys.exit(main())
