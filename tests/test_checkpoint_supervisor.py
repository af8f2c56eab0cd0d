"""CPU tests: Saver state file/rotation, FileWriter, Supervisor lifecycle with an in-memory client."""
import os
import threading
import time

import numpy as np
import pytest

from distributedtensorflowexample_b200.parallel.client import ParamClient
from distributedtensorflowexample_b200.train.supervisor import Supervisor
from distributedtensorflowexample_b200.utils import checkpoint as ckpt
from distributedtensorflowexample_b200.utils.events import FileWriter, training_summary, TAG_LOSS

NAMES = ["global/dense/kernel", "global/dense/bias", "global/dense_1/kernel", "global/dense_1/bias", "global/global_step"]


def _values(step):
    return {NAMES[0]: np.full((784, 100), step, np.float32), NAMES[1]: np.zeros(100, np.float32),
            NAMES[2]: np.ones((100, 10), np.float32), NAMES[3]: np.zeros(10, np.float32),
            NAMES[4]: np.array(step, np.int32)}


def test_saver_files_state_and_rotation(tmp_path):
    d = str(tmp_path)
    s = ckpt.FastSaver(NAMES)
    paths = [s.save(_values(k), os.path.join(d, "model.ckpt"), global_step=k * 10) for k in range(7)]
    assert paths[-1].endswith("model.ckpt-60")
    files = sorted(os.listdir(d))
    assert "checkpoint" in files and not any(f.endswith(".meta") for f in files)
    kept = [f"model.ckpt-{k * 10}" for k in range(2, 7)]                      # max_to_keep = 5
    assert sorted(f for f in files if f.endswith(".index")) == sorted(k + ".index" for k in kept)
    text = open(os.path.join(d, "checkpoint")).read().splitlines()
    assert text[0] == f'model_checkpoint_path: "{os.path.join(d, "model.ckpt-60")}"'
    assert text[1:] == [f'all_model_checkpoint_paths: "{os.path.join(d, k)}"' for k in kept]    # oldest first
    assert ckpt.latest_checkpoint(d) == os.path.join(d, "model.ckpt-60")
    back = ckpt.FastSaver(NAMES).restore(ckpt.latest_checkpoint(d))
    assert back[NAMES[4]].dtype == np.int32 and int(back[NAMES[4]]) == 6 and back[NAMES[0]][0, 0] == 6.0
    # a new Saver in a restarted process continues the rotation from the state file
    s2 = ckpt.FastSaver(NAMES)
    s2.save(_values(7), os.path.join(d, "model.ckpt"), global_step=70)
    assert not os.path.exists(os.path.join(d, "model.ckpt-20.index"))
    assert ckpt.latest_checkpoint(str(tmp_path / "nothing")) is None


class MemClient(ParamClient):
    def __init__(self):
        self.values, self.ready, self.inits = {}, False, 0
        self.lock = threading.Lock()

    def initialize(self, rng):
        self.inits += 1
        self.values = _values(0)

    def assign_many(self, values):
        self.values = {k: np.array(v) for k, v in values.items()}

    def snapshot(self, names=None):
        with self.lock:
            return {n: np.array(self.values[n]) for n in (names or self.values)}

    def read_step(self):
        return int(self.values[NAMES[4]])

    def bump(self, k=1):
        with self.lock:
            self.values[NAMES[4]] = np.array(self.read_step() + k, np.int32)

    def mark_ready(self):
        self.ready = True

    def report_uninitialized_variables(self):
        return [] if self.ready else list(NAMES)


def _sv(client, logdir, chief=True, writer=None, **kw):
    return Supervisor(is_chief=chief, logdir=logdir, saver=ckpt.FastSaver(NAMES), client=client, summary_writer=writer,
                      init_op=lambda: client.initialize(None), graph_text="node {}\n", **kw)


def test_supervisor_chief_init_checkpoint_then_resume(tmp_path):
    logdir = str(tmp_path / "log")
    c = MemClient()
    w = FileWriter(logdir + "_0", flush_secs=0.2)
    sv = _sv(c, logdir, writer=w, save_model_secs=0.2, save_summaries_secs=0.1)
    with sv.managed_session() as sess:
        assert sess is c and c.inits == 1 and c.ready and sv.restored_from is None
        assert os.path.exists(os.path.join(logdir, "graph.pbtxt"))
        time.sleep(0.15)
        assert os.path.exists(os.path.join(logdir, "model.ckpt-0.index"))         # first save fires immediately
        c.bump(42)
        time.sleep(0.5)
        assert not sv.should_stop()
    assert sv.should_stop()
    assert ckpt.latest_checkpoint(logdir).endswith("model.ckpt-42")
    w.close()
    # restart: the chief restores weights AND global_step instead of re-initialising
    c2 = MemClient()
    sv2 = _sv(c2, logdir, save_model_secs=0, save_summaries_secs=0)
    with sv2.managed_session():
        assert c2.inits == 0 and c2.read_step() == 42 and sv2.restored_from.endswith("model.ckpt-42")
    from tensorboard.backend.event_processing.event_file_loader import LegacyEventFileLoader
    evs = list(LegacyEventFileLoader(w.path).Load())
    assert evs[0].file_version == "brain.Event:2"
    assert any(e.HasField("graph_def") for e in evs)
    assert [e.session_log.status for e in evs if e.HasField("session_log")][0] == 1            # START
    assert any(e.session_log.checkpoint_path.endswith("model.ckpt-42") for e in evs if e.HasField("session_log"))
    assert any(v.tag == "global/global_step/sec" for e in evs for v in e.summary.value)


def test_supervisor_non_chief_waits_for_chief(tmp_path):
    c = MemClient()
    sv = _sv(c, str(tmp_path / "l"), chief=False, recovery_wait_secs=0.05)
    entered = []

    def run():
        with sv.managed_session():
            entered.append(time.time())
    t = threading.Thread(target=run); t.start()
    time.sleep(0.2)
    assert not entered and c.inits == 0            # still polling ready_op; a non-chief never initialises
    c.initialize(None); c.mark_ready()
    t.join(timeout=5)
    assert entered and not os.path.exists(str(tmp_path / "l" / "graph.pbtxt"))     # only the chief writes


def test_supervisor_stops_on_exception_and_service_failure(tmp_path):
    c = MemClient()
    sv = _sv(c, str(tmp_path / "a"), save_model_secs=0.05)
    with pytest.raises(SystemExit) as ei:
        with sv.managed_session():
            raise SystemExit(143)                  # what the SIGTERM handler raises (reference main.py:51-52)
    assert ei.value.code == 143 and sv.should_stop()
    c2 = MemClient()
    sv2 = _sv(c2, str(tmp_path / "b"), save_model_secs=0.05)
    with sv2.managed_session():
        c2.values.pop(NAMES[0])                    # next checkpoint attempt fails -> coordinator stops training
        for _ in range(100):
            if sv2.should_stop():
                break
            time.sleep(0.02)
        assert sv2.should_stop()


def test_file_writer_per_step_summaries(tmp_path):
    w = FileWriter(str(tmp_path / "run_3"))
    assert os.path.basename(w.path).startswith("events.out.tfevents.")
    for s in range(10):
        w.add_summary(training_summary(1.0 / (s + 1), s / 10.0), s)
    w.close()
    from tensorboard.backend.event_processing.event_file_loader import LegacyEventFileLoader
    evs = [e for e in LegacyEventFileLoader(w.path).Load() if e.summary.value]
    assert [e.step for e in evs] == list(range(10))
    assert [v.tag for v in evs[0].summary.value] == [TAG_LOSS, "local/accuracy_1"]
