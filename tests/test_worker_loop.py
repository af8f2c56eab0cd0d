"""The GPU worker loop's host logic, without a GPU: launch-size planning (steps_per_launch adaptive) and the
per-launch digest that applies the reference's cadence rules (worker.py:139-159) to all steps of a launch."""
import types

import numpy as np
from hypothesis import given, settings, strategies as st

from distributedtensorflowexample_b200.train import worker as W
from distributedtensorflowexample_b200.utils import tracing


def _simulate_single_worker(T, E, cap, start=0):
    lb, launches = start, []
    while True:
        n = W.plan_launch_steps(lb, T, E, 1, cap, True)
        assert 1 <= n <= cap
        launches.append((lb, n))
        lb += n
        if lb - 1 >= T:
            return launches


@settings(max_examples=200, deadline=None)
@given(T=st.integers(0, 30000), E=st.sampled_from([0, 7, 100, 10000]), cap=st.integers(1, 600), start=st.integers(0, 20000))
def test_single_worker_runs_exactly_the_reference_step_count(T, E, cap, start):
    """A lone worker observes steps start, start+1, ... and the reference stops after the first step that observes
    >= training_steps: T - start + 1 steps (one step if it starts beyond T)."""
    launches = _simulate_single_worker(T, E, cap, start)
    assert sum(n for _, n in launches) == max(1, T - start + 1)
    if E:
        for lb, n in launches:      # an evaluation step is never in the middle of a launch
            inside = [s for s in range(lb, lb + n - 1) if s % E == 0]
            assert not inside, (lb, n, inside)


@settings(max_examples=200, deadline=None)
@given(lb=st.integers(0, 10 ** 7), T=st.integers(0, 10 ** 7), Wk=st.integers(1, 16), cap=st.integers(1, 512))
def test_plan_is_bounded_and_shrinks_towards_the_end(lb, T, Wk, cap):
    n = W.plan_launch_steps(lb, T, 10000, Wk, cap, False)
    assert 0 <= n <= cap
    if lb > T:
        assert n == 0 and W.plan_launch_steps(lb, T, 10000, Wk, cap, True) == 1
    else:
        assert n >= 1 and (n - 1) * Wk <= T - lb        # even if every other worker keeps pace, no overshoot


class _Rec:
    def __init__(self): self.items = []
    def submit(self, s, l, a): self.items.append((s.copy(), l.copy(), a.copy()))


def _worker(training_steps=10 ** 7, log_every=100, eval_every=10000, faults=None):
    w = object.__new__(W.Worker)
    w.flags = types.SimpleNamespace(training_steps=training_steps, log_every=log_every, eval_every=eval_every)
    w.faults = faults or {}
    w.task_index = 0
    w.tracer = tracing.get_tracer()
    w._start_time, w._start_step = 100.0, 0
    w.summary_writer = types.SimpleNamespace(added=[], add_summary=lambda summ, step: w.summary_writer.added.append((step, summ.values)))
    return w


def test_digest_applies_the_cadence_rules_to_every_step_of_a_launch(capsys):
    w = _worker(training_steps=450)
    rec = _Rec()
    steps = np.arange(150, 350, dtype=np.int32)           # one 200-step launch, a lone worker
    loss = np.linspace(2.0, 1.0, 200, dtype=np.float32)
    acc = np.full(200, 0.5, np.float32)
    evals = []
    last, stop = w._digest_launch(loss, acc, steps, 101.0, 103.0, lambda: evals.append(1) or 0.9, rec)
    out = capsys.readouterr().out.splitlines()
    assert (last, stop) == (349, False) and not evals
    assert [l.split("\t")[0] for l in out] == ["step: 200", "step: 300"]
    # speeds use interpolated step times: step 200 is the 51st of 200 steps spread over [101, 103]
    sp = [float(l.split("speed: ")[1].replace("step/sec", "")) for l in out]
    assert abs(sp[0] - 200 / (101.0 + 2.0 * 51 / 200 - 100.0)) < 1e-6
    assert abs(sp[1] - 100 / (2.0 * 100 / 200)) < 1e-6
    assert len(rec.items) == 1 and rec.items[0][0].tolist() == list(range(150, 350))
    # the launch that crosses training_steps: summaries and rules stop at the first step >= training_steps
    steps = np.arange(350, 550, dtype=np.int32)
    last, stop = w._digest_launch(loss, acc, steps, 103.0, 105.0, lambda: 0.9, rec)
    assert (last, stop) == (450, True)
    assert rec.items[1][0].tolist() == list(range(350, 451))


def test_digest_evaluates_at_multiples_of_eval_every_including_zero(capsys):
    w = _worker(eval_every=50, log_every=1000)
    rec = _Rec()
    steps = np.array([0], dtype=np.int32)
    w._digest_launch(np.ones(1, np.float32), np.ones(1, np.float32), steps, 0.0, 1.0, lambda: 0.25, rec)
    steps = np.arange(1, 101, dtype=np.int32) * 2        # another worker takes every other step
    w._digest_launch(np.ones(100, np.float32), np.ones(100, np.float32), steps, 1.0, 2.0, lambda: 0.5, rec)
    out = capsys.readouterr().out.splitlines()
    assert out == ["test accuracy: 0.25"] + ["test accuracy: 0.5"] * 4          # steps 50, 100, 150, 200
    # every evaluation is also logged (additive tag), keyed by the step that triggered it
    assert [s for s, _ in w.summary_writer.added] == [0, 50, 100, 150, 200]
    assert w.summary_writer.added[1][1] == [("local/test_accuracy", 0.5)]


def _race(T, Wk, cap, E, seed, depth=W.PIPELINE_DEPTH):
    """Wk workers with the loop's planning rule and two launches in flight each, steps interleaved at random."""
    import random
    rnd = random.Random(seed)
    g = 0
    ws = [dict(pending=[], next_lb=0, inflight=0, stop=False, hold=False) for _ in range(Wk)]
    while not all(w["stop"] for w in ws):
        w = rnd.choice([w for w in ws if not w["stop"]])
        while len(w["pending"]) < depth and not w["hold"]:
            lb = w["next_lb"] + Wk * w["inflight"]
            n = W.plan_launch_steps(lb, T, E, Wk, cap, not w["pending"])
            if n == 0:
                break
            w["hold"] = E > 0 and lb + n - 1 >= -(-lb // E) * E
            w["pending"].append([n, []]); w["inflight"] += n
        l = w["pending"][0]
        l[1].append(g); g += 1
        if len(l[1]) == l[0]:
            w["pending"].pop(0); w["inflight"] -= l[0]
            hit = [s for s in l[1] if s >= T]
            w["stop"] = bool(hit)
            w["next_lb"] = max(w["next_lb"], (hit[0] if hit else l[1][-1]) + 1)
            if not w["pending"]:
                w["hold"] = False
    return g


def test_racing_workers_overshoot_training_steps_by_little():
    """The reference overshoots training_steps by at most num_workers - 1 updates (every worker notices at its next
    step).  Multi-step launches notice a launch late; the planning rule keeps that to a few dozen updates per worker
    instead of the thousands a fixed launch size would run past the bound."""
    for Wk in (2, 4, 8, 16):
        for T in (250, 5000, 60000):
            over = max(_race(T, Wk, 256, 10000, seed) - T for seed in range(3))
            assert 0 < over <= 24 * Wk, (Wk, T, over)
    assert _race(5000, 1, 256, 10000, 0) == 5001          # a lone worker: exactly the reference's count
