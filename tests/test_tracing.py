"""Tracing subsystem: spans, counters, Chrome-trace output, disabled-by-default cost."""
import json
import threading
import time

from distributedtensorflowexample_b200.utils import tracing


def test_disabled_tracer_is_a_noop(tmp_path):
    tr = tracing.Tracer(False)
    with tr.span("x"):
        pass
    with tr.device_span("y"):
        pass
    tr.instant("i"); tr.counter("c", v=1)
    assert tr.events() == [] and tr.save(str(tmp_path / "t.json")) is None
    assert tr.span("a") is tr.span("b")                 # the shared null context: nothing is allocated per span


def test_spans_counters_and_chrome_trace_file(tmp_path):
    tr = tracing.Tracer(True, process_name="worker:3", nvtx=False)
    with tr.span("outer", "step", steps=4):
        time.sleep(0.002)
        with tr.span("inner"):
            time.sleep(0.001)
    tr.instant("checkpoint", step=7)
    tr.counter("speed", steps_per_sec=123.0)

    def other():
        with tr.span("thread_span"):
            pass
    t = threading.Thread(target=other); t.start(); t.join()

    path = tr.save(str(tmp_path / "sub" / "trace.json"))
    doc = json.load(open(path))
    evs = doc["traceEvents"]
    by = {e["name"]: e for e in evs if e.get("ph") in ("X", "i", "C")}
    assert by["outer"]["args"] == {"steps": 4} and by["outer"]["cat"] == "step"
    assert by["outer"]["dur"] >= by["inner"]["dur"] >= 900            # microseconds
    assert by["outer"]["ts"] <= by["inner"]["ts"]
    assert by["inner"]["ts"] + by["inner"]["dur"] <= by["outer"]["ts"] + by["outer"]["dur"] + 1
    assert by["thread_span"]["tid"] != by["outer"]["tid"]
    assert by["checkpoint"]["ph"] == "i" and by["speed"]["args"] == {"steps_per_sec": 123.0}
    assert any(e.get("ph") == "M" and e["args"].get("name") == "worker:3" for e in evs)
    s = tr.summary()
    assert s["outer"]["count"] == 1 and s["inner"]["mean_us"] >= 900


def test_event_cap_and_task_paths(tmp_path):
    tr = tracing.Tracer(True, max_events=3, nvtx=False)
    for i in range(10):
        tr.instant(f"e{i}")
    assert len(tr.events()) == 3 and tr.dropped == 7
    assert tracing.trace_path_for_task("/x/run.json", "worker", 2) == "/x/run.worker2.json"
    assert tracing.trace_path_for_task("run", "ps", 0) == "run.ps0.json"


def test_global_enable_disable():
    tracing.disable()
    assert not tracing.get_tracer().enabled
    a = tracing.enable("p")
    assert a.enabled and tracing.enable("q") is a
    tracing.disable()
    assert not tracing.get_tracer().enabled
