"""GPU: the reference's process model on real hardware - a ps process owning an HBM shard plus worker
processes that map it over CUDA IPC (run_single_gpu.sh topology: 2 workers sharing GPU 0)."""
import re
import socket

import numpy as np
import pytest

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(600)]

from distributedtensorflowexample_b200.launch import launcher
from distributedtensorflowexample_b200.utils import checkpoint as ckpt


def _port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _run(tmp_path, extra, num_workers=2, num_gpus=1, gpu_ids=("0",), num_ps=1, timeout=300):
    logdir = str(tmp_path / "log")
    args = ["--logdir", logdir, "--seed", "1", "--recovery_wait_secs", "0.5", "--save_model_secs", "0.3",
            "--base_port", str(_port()), "--data", "synthetic", *extra]
    tasks = launcher.task_commands(num_workers, num_gpus, list(gpu_ids), num_ps=num_ps, device="cuda", extra=args)
    cl = launcher.HeadlessCluster(tasks, echo=False)
    cl.start()
    codes = cl.wait(timeout=timeout)
    return cl, codes, logdir


def test_async_two_workers_one_gpu_then_resume(tmp_path):
    cl, codes, logdir = _run(tmp_path, ["--training_steps", "3000", "--learning_rate", "0.05"])
    assert codes["worker0"] == 0 and codes["worker1"] == 0, (codes, cl.lines)
    assert abs(codes["ps"]) in (15, 143)                       # the ps only ever leaves by signal, like the reference
    w0 = cl.lines["worker0"]
    assert any(l.startswith("Per-process GPU memory fraction: 0.45") for l in w0)
    # printed by whichever worker observes global step 0 (usually the chief, but the other worker may win the race)
    assert any(l.startswith("test accuracy: ") for l in w0 + cl.lines["worker1"])
    pat = re.compile(r"^step: (\d+)\t\| cost: ([-0-9.e]+)\t\| speed: ([-0-9.e]+)step/sec$")
    costs = [float(m.group(2)) for l in w0 + cl.lines["worker1"] if (m := pat.match(l))]
    assert costs and min(costs) < 2.0 < max(costs) + 20         # it trains from ~9 down
    latest = ckpt.latest_checkpoint(logdir)
    assert latest is not None
    vals = ckpt.FastSaver(["global/dense/kernel", "global/global_step"]).restore(latest)
    assert vals["global/dense/kernel"].shape == (784, 100) and vals["global/global_step"].dtype == np.int32
    step_a = int(vals["global/global_step"])
    cl2, codes2, _ = _run(tmp_path, ["--training_steps", str(step_a + 2000), "--learning_rate", "0.05"])
    assert codes2["worker0"] == 0, cl2.lines
    assert any(l.startswith("restored from ") for l in cl2.lines["worker0"])


def test_adam_and_unlocked_apply_modes(tmp_path):
    cl, codes, _ = _run(tmp_path, ["--training_steps", "600", "--optimizer", "adam", "--learning_rate", "0.01"])
    assert codes["worker0"] == 0 and codes["worker1"] == 0, (codes, cl.lines)
    cl, codes, _ = _run(tmp_path / "u", ["--training_steps", "600", "--unlocked_apply", "--steps_per_launch", "10"])
    assert codes["worker0"] == 0 and codes["worker1"] == 0, (codes, cl.lines)


def test_bfloat16_worker_trains_on_the_tensor_core_engine(tmp_path):
    """--dtype bfloat16 selects the tcgen05 engine: data set, test set and parameter handles must all follow."""
    cl, codes, logdir = _run(tmp_path, ["--training_steps", "300", "--dtype", "bfloat16", "--hidden", "104",
                                        "--learning_rate", "0.05", "--eval_every", "200"], num_workers=1)
    assert codes["worker0"] == 0, (codes, cl.lines)
    assert sum(l.startswith("test accuracy: ") for l in cl.lines["worker0"]) >= 2        # steps 0 and 200


def test_default_flags_reach_training_steps_exactly_with_one_worker(tmp_path):
    """Adaptive launch sizes: a lone worker runs exactly training_steps + 1 steps (the reference's count) and the
    evaluation steps end their launches."""
    cl, codes, logdir = _run(tmp_path, ["--training_steps", "2500", "--learning_rate", "0.05", "--eval_every", "1000"],
                             num_workers=1)
    assert codes["worker0"] == 0, (codes, cl.lines)
    w0 = cl.lines["worker0"]
    assert sum(l.startswith("test accuracy: ") for l in w0) == 3                          # steps 0, 1000, 2000
    pat = re.compile(r"^step: (\d+)\t")
    assert [int(m.group(1)) for l in w0 if (m := pat.match(l))] == list(range(100, 2501, 100))
    from distributedtensorflowexample_b200.utils.events_reader import scalars
    assert [s for s, _, _, _ in scalars(logdir + "_0", "local/loss")] == list(range(2501))


@pytest.mark.multigpu
def test_sync_mode_two_gpus_two_ps(tmp_path):
    cl, codes, logdir = _run(tmp_path, ["--training_steps", "400", "--sync", "--learning_rate", "0.05"],
                             num_workers=2, num_gpus=2, gpu_ids=("0", "1"), num_ps=2)
    assert codes["worker0"] == 0 and codes["worker1"] == 0, (codes, cl.lines)
    # both workers observe the same aggregated-update counter, so they log the same steps
    pat = re.compile(r"^step: (\d+)\t")
    s0 = [int(m.group(1)) for l in cl.lines["worker0"] if (m := pat.match(l))]
    s1 = [int(m.group(1)) for l in cl.lines["worker1"] if (m := pat.match(l))]
    assert s0 and s0 == s1
