"""Event-file reader and the two command-line tools (inspect_checkpoint, read_events)."""
import os
import subprocess
import sys

import numpy as np
import pytest

from distributedtensorflowexample_b200.tools import inspect_checkpoint, read_events
from distributedtensorflowexample_b200.utils import checkpoint as ckpt
from distributedtensorflowexample_b200.utils import events_reader, native_io as nio

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _write_events(path):
    w = nio.EventsFile(path)
    w.write_version(100.0)
    w.write_session_log(100.5, 0, nio.SESSION_START)
    w.write_graph(100.6, b"node { name: 'x' }")
    for s in range(5):
        w.write_scalars(101.0 + s, s, ["local/loss", "local/accuracy_1"], [9.75 - s, 0.1 * s])
    w.write_session_log(110.0, 4, nio.SESSION_CHECKPOINT, checkpoint_path="/x/model.ckpt-4")
    w.write_scalars(111.0, 4, ["global/global_step/sec"], [12345.5])
    w.flush(); w.close()


def test_events_reader_decodes_what_the_writer_wrote(tmp_path):
    path = str(tmp_path / "events.out.tfevents.0000000123.host")
    _write_events(path)
    evs = list(events_reader.read_events(path))
    assert len(evs) == 10 and evs[0].file_version == "brain.Event:2" and evs[0].wall_time == 100.0
    assert evs[1].session_log == {"status": 1, "checkpoint_path": "", "msg": ""}
    assert evs[2].graph_def_bytes == len(b"node { name: 'x' }")
    assert evs[3].step == 0 and evs[6].step == 3
    assert dict(evs[6].scalars)["local/loss"] == pytest.approx(6.75) and dict(evs[6].scalars)["local/accuracy_1"] == pytest.approx(0.3)
    assert evs[8].session_log["status"] == 3 and evs[8].session_log["checkpoint_path"] == "/x/model.ckpt-4"
    sc = events_reader.scalars(str(tmp_path), tag="local/loss")                 # directory form
    assert [s for s, _, _, _ in sc] == [0, 1, 2, 3, 4]
    # same view as tensorboard's own loader, when it is installed
    loader = pytest.importorskip("tensorboard.backend.event_processing.event_file_loader")
    ref = list(loader.LegacyEventFileLoader(path).Load())
    assert [e.step for e in ref] == [e.step for e in evs]
    assert [[(v.tag, v.simple_value) for v in e.summary.value] for e in ref] == [e.scalars for e in evs]


def test_events_reader_detects_corruption_and_tolerates_a_truncated_tail(tmp_path):
    path = str(tmp_path / "events.out.tfevents.1.h")
    _write_events(path)
    raw = open(path, "rb").read()
    open(path, "wb").write(raw[:-7])                                    # a writer caught mid-record
    assert len(list(events_reader.read_events(path))) == 9
    bad = bytearray(raw); bad[40] ^= 0x55
    open(path, "wb").write(bytes(bad))
    with pytest.raises(events_reader.EventFileError):
        list(events_reader.read_events(path))


def test_inspect_checkpoint_and_read_events_tools(tmp_path, capsys):
    logdir = str(tmp_path / "log")
    tensors = {"global/dense/kernel": np.arange(12, dtype=np.float32).reshape(3, 4),
               "global/global_step": np.array(42, dtype=np.int32)}
    saver = ckpt.FastSaver(list(tensors))
    prefix = saver.save(tensors, os.path.join(logdir, "model.ckpt"), global_step=42)
    text = inspect_checkpoint.describe(inspect_checkpoint.resolve_prefix(logdir))
    assert "global/dense/kernel (DT_FLOAT32) [3, 4]" in text and "global/global_step (DT_INT32) []" in text
    assert "# 2 tensors, 13 elements" in text
    assert inspect_checkpoint.resolve_prefix(prefix + ".index") == prefix
    assert "11." in inspect_checkpoint.describe(prefix, tensor="global/dense/kernel")
    assert inspect_checkpoint.main([str(tmp_path / "nothing")]) == 1
    os.makedirs(logdir + "_0", exist_ok=True)
    _write_events(os.path.join(logdir + "_0", "events.out.tfevents.5.h"))
    out = read_events.format_events(logdir + "_0", tag="local/loss")
    assert out.splitlines()[0].endswith("step 0 local/loss=9.75") and len(out.splitlines()) == 5
    full = read_events.format_events(logdir + "_0")
    assert "session_log CHECKPOINT /x/model.ckpt-4" in full and "file_version brain.Event:2" in full
    # the module entry points work as commands
    r = subprocess.run([sys.executable, "-m", "distributedtensorflowexample_b200.tools.inspect_checkpoint", logdir],
                       cwd=ROOT, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and "global/global_step" in r.stdout
    r = subprocess.run([sys.executable, "-m", "distributedtensorflowexample_b200.tools.read_events", logdir + "_0",
                        "--limit", "2"], cwd=ROOT, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and len(r.stdout.strip().splitlines()) == 2
