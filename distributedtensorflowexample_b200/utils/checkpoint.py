"""Saver with TF1 semantics on top of the native tensor-bundle writer (reference utils.py:28-32
``FastSaver``, worker.py:102-103; TF1 tf.train.Saver - SURVEY C15/T13, Appendix B).

Files: ``<save_path>-<global_step>.index`` + ``.data-00000-of-00001`` (no ``.meta``: the reference's
FastSaver forces ``write_meta_graph=False``), and the text-proto state file ``<dir>/checkpoint``
listing the kept checkpoints oldest first.  ``max_to_keep=5`` like TF1's default.
"""
from __future__ import annotations

import os
import re
from typing import Dict, List, Optional, Sequence

import numpy as np

from . import native_io

STATE_FILE = "checkpoint"
_LINE = re.compile(r'^\s*(model_checkpoint_path|all_model_checkpoint_paths)\s*:\s*"(.*)"\s*$')


def _escape(p: str) -> str:
    return p.replace("\\", "\\\\").replace('"', '\\"')


def _unescape(p: str) -> str:
    return p.replace('\\"', '"').replace("\\\\", "\\")


class CheckpointState:
    def __init__(self, model_checkpoint_path: str = "", all_model_checkpoint_paths: Optional[List[str]] = None):
        self.model_checkpoint_path = model_checkpoint_path
        self.all_model_checkpoint_paths = list(all_model_checkpoint_paths or [])

    def to_text(self) -> str:
        lines = [f'model_checkpoint_path: "{_escape(self.model_checkpoint_path)}"']
        lines += [f'all_model_checkpoint_paths: "{_escape(p)}"' for p in self.all_model_checkpoint_paths]
        return "\n".join(lines) + "\n"

    @staticmethod
    def from_text(text: str) -> "CheckpointState":
        st = CheckpointState()
        for line in text.splitlines():
            m = _LINE.match(line)
            if not m:
                continue
            if m.group(1) == "model_checkpoint_path":
                st.model_checkpoint_path = _unescape(m.group(2))
            else:
                st.all_model_checkpoint_paths.append(_unescape(m.group(2)))
        return st


def get_checkpoint_state(checkpoint_dir: str, latest_filename: str = STATE_FILE) -> Optional[CheckpointState]:
    path = os.path.join(checkpoint_dir, latest_filename)
    if not os.path.exists(path):
        return None
    with open(path, "r") as f:
        st = CheckpointState.from_text(f.read())
    # relative entries are relative to the checkpoint directory (TF1 behaviour)
    fix = lambda p: p if os.path.isabs(p) or os.path.exists(p + ".index") else os.path.join(checkpoint_dir, p)
    st.model_checkpoint_path = fix(st.model_checkpoint_path) if st.model_checkpoint_path else ""
    st.all_model_checkpoint_paths = [fix(p) for p in st.all_model_checkpoint_paths]
    return st


def checkpoint_exists(prefix: str) -> bool:
    return os.path.exists(prefix + ".index") and os.path.exists(prefix + ".data-00000-of-00001")


def latest_checkpoint(checkpoint_dir: str, latest_filename: str = STATE_FILE) -> Optional[str]:
    st = get_checkpoint_state(checkpoint_dir, latest_filename)
    if st and st.model_checkpoint_path and checkpoint_exists(st.model_checkpoint_path):
        return st.model_checkpoint_path
    return None


class Saver:
    """Saves/restores a fixed list of named variables.

    ``var_list``: names to save (the reference passes ``get_vars('global', False)`` = 4 weights +
    ``global/global_step``).  ``save(values, save_path, global_step)`` takes the current values as
    numpy arrays (a fuzzy snapshot of the shards)."""

    def __init__(self, var_list: Sequence[str], max_to_keep: int = 5):
        self.var_list = list(var_list)
        self.max_to_keep = max_to_keep
        self._kept: List[str] = []

    def save(self, values: Dict[str, np.ndarray], save_path: str, global_step: Optional[int] = None,
             latest_filename: str = STATE_FILE, write_meta_graph: bool = True) -> str:
        missing = [n for n in self.var_list if n not in values]
        if missing:
            raise KeyError(f"Saver.save: no value for {missing}")
        prefix = save_path if global_step is None else f"{save_path}-{int(global_step)}"
        d = os.path.dirname(prefix) or "."
        os.makedirs(d, exist_ok=True)
        native_io.bundle_write(prefix, {n: values[n] for n in self.var_list})
        if write_meta_graph:
            # A TF1 .meta holds the GraphDef; there is no graph here.  Kept as an explicit no-op so that
            # FastSaver's "never write it" is a visible decision rather than an accident.
            pass
        self._record(prefix, d, latest_filename)
        return prefix

    def _record(self, prefix: str, d: str, latest_filename: str) -> None:
        if not self._kept:
            st = get_checkpoint_state(d, latest_filename)
            if st:
                self._kept = [p for p in st.all_model_checkpoint_paths if checkpoint_exists(p)]
        if prefix in self._kept:
            self._kept.remove(prefix)
        self._kept.append(prefix)
        while self.max_to_keep and len(self._kept) > self.max_to_keep:
            old = self._kept.pop(0)
            for suffix in (".index", ".data-00000-of-00001", ".meta"):
                try:
                    os.remove(old + suffix)
                except FileNotFoundError:
                    pass
        state = CheckpointState(prefix, self._kept)
        tmp = os.path.join(d, latest_filename + f".tmp{os.getpid()}")
        with open(tmp, "w") as f:
            f.write(state.to_text())
        os.replace(tmp, os.path.join(d, latest_filename))

    def restore(self, prefix: str) -> Dict[str, np.ndarray]:
        vals = native_io.bundle_read(prefix)
        missing = [n for n in self.var_list if n not in vals]
        if missing:
            raise KeyError(f"checkpoint {prefix} lacks {missing}")
        return {n: vals[n] for n in self.var_list}

    @property
    def last_checkpoints(self) -> List[str]:
        return list(self._kept)


class FastSaver(Saver):
    """Reference ``FastSaver`` (utils.py:28-32): a Saver whose ``save`` never writes the meta graph."""

    def save(self, values, save_path, global_step=None, latest_filename=STATE_FILE, meta_graph_suffix="meta",
             write_meta_graph=True):
        return super().save(values, save_path, global_step, latest_filename, write_meta_graph=False)
