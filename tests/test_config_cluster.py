"""CPU unit tests: flags, cluster spec, placement, variable scopes, memory fraction, launchers."""
import os
import subprocess
import sys

import pytest

from distributedtensorflowexample_b200 import flags as F
from distributedtensorflowexample_b200.cluster import ClusterSpec, cluster_spec
from distributedtensorflowexample_b200.placement import (DeviceSpec, GreedyLoadBalancingStrategy, OpInfo,
                                                         replica_device_setter)
from distributedtensorflowexample_b200.utils import memory
from distributedtensorflowexample_b200.variables import VariableStore, get_vars, random_normal_initializer


def _flags():
    fv = F.FlagValues()
    F.define_reference_flags(fv)
    F.define_framework_flags(fv)
    return fv


def test_reference_flag_defaults():
    fv = _flags()
    fv.parse([])
    assert (fv.job_name, fv.task_index, fv.batch_size, fv.learning_rate) == ("ps", 0, 100, 0.001)
    assert (fv.training_steps, fv.logdir, fv.num_workers, fv.num_gpus) == (10 ** 7, "./tmp/mnist/1", 2, 1)
    assert fv.num_ps == 1 and fv.sync is False and fv.optimizer == "sgd"


def test_flag_spellings_and_unknown_tolerated():
    fv = _flags()
    rest = fv.parse(["--num_workers", "16", "--num_gpus=4", "--job_name", "worker", "--task_index=3",
                     "--totally_unknown=1", "--learning_rate", "0.1", "--sync", "--training_steps=1e4", "positional"])
    assert (fv.num_workers, fv.num_gpus, fv.job_name, fv.task_index) == (16, 4, "worker", 3)
    assert fv.learning_rate == 0.1 and fv.sync is True and fv.training_steps == 10000
    assert rest == ["--totally_unknown=1", "positional"] and fv.unparsed == rest


def test_flags_lazy_parse(monkeypatch):
    fv = _flags()
    monkeypatch.setattr(sys, "argv", ["prog", "--batch_size=7"])
    assert fv.batch_size == 7              # parsed on first access, no explicit parse()/app.run()
    fv.batch_size = 9
    assert fv.batch_size == 9
    with pytest.raises(AttributeError):
        fv.nope
    with pytest.raises(ValueError):
        _flags().parse(["--task_index", "abc"])
    fv2 = _flags(); fv2.parse(["--nosync", "--unlocked_apply=true"])
    assert fv2.sync is False and fv2.unlocked_apply is True


def test_cluster_spec_ports_and_order():
    spec = cluster_spec(2, 1)
    assert spec == {"ps": ["127.0.0.1:12222"], "worker": ["127.0.0.1:12223", "127.0.0.1:12224"]}
    spec = cluster_spec(16, 3)
    assert spec["ps"] == [f"127.0.0.1:{12222 + i}" for i in range(3)]
    assert spec["worker"][0] == "127.0.0.1:12225" and spec["worker"][-1] == "127.0.0.1:12240"
    cs = ClusterSpec(spec)
    assert cs.num_tasks("worker") == 16 and cs.task_address("ps", 2) == "127.0.0.1:12224"
    assert cs.task_host_port("worker", 1) == ("127.0.0.1", 12226)
    assert cs.global_rank("ps", 1) == 1 and cs.global_rank("worker", 0) == 3 and cs.world_size == 19
    with pytest.raises(ValueError):
        cs.task_address("worker", 16)
    with pytest.raises(ValueError):
        ClusterSpec({"ps": ["nohostport"]})


def test_device_spec_roundtrip_and_merge():
    d = DeviceSpec.from_string("/job:worker/task:3/gpu:0")
    assert (d.job, d.task, d.device_type, d.device_index) == ("worker", 3, "gpu", 0)
    assert d.to_string() == "/job:worker/task:3/gpu:0"
    assert DeviceSpec.from_string("/job:ps").merge_from(DeviceSpec.from_string("/task:1/cpu:0")).to_string() == "/job:ps/task:1/cpu:0"
    with pytest.raises(ValueError):
        DeviceSpec.from_string("garbage")


@pytest.mark.parametrize("num_ps", [1, 2, 3])
def test_replica_device_setter_round_robin(num_ps):
    fn = replica_device_setter(num_ps, worker_device="/job:worker/task:0/cpu:0")
    devs = [fn(OpInfo(f"v{i}", True, 4)) for i in range(5)]
    assert devs == [f"/job:ps/task:{i % num_ps}" for i in range(5)]
    assert fn(OpInfo("matmul", False)) == "/job:worker/task:0/cpu:0"
    assert replica_device_setter(0, worker_device="/job:worker/task:0")(OpInfo("v", True)) == "/job:worker/task:0"


def test_greedy_strategy_balances_bytes():
    fn = replica_device_setter(2, ps_strategy=GreedyLoadBalancingStrategy(2))
    sizes = [313600, 400, 4000, 40, 4]
    devs = [fn(OpInfo(f"v{i}", True, s)) for i, s in enumerate(sizes)]
    assert devs[0] == "/job:ps/task:0" and all(d == "/job:ps/task:1" for d in devs[1:])


def test_variable_scopes_names_and_get_vars():
    from distributedtensorflowexample_b200.models.mlp import reference_mlp_spec
    from distributedtensorflowexample_b200.train.replica import build_replicated_model
    m = build_replicated_model(reference_mlp_spec(), "worker", 1, num_ps=2)
    g = [v.name for v in m.global_trainable]
    assert g == ["global/dense/kernel", "global/dense/bias", "global/dense_1/kernel", "global/dense_1/bias"]
    assert [v.name for v in m.local_trainable] == [n.replace("global", "local") for n in g]
    assert [v.name for v in m.global_all] == g + ["global/global_step"]         # the saver's 5 tensors
    assert m.global_step.dtype == "int32" and m.global_step.shape == () and not m.global_step.trainable
    assert [v.shape for v in m.global_trainable] == [(784, 100), (100,), (100, 10), (10,)]
    assert sum(v.numel for v in m.global_trainable) == 79510
    assert [v.device.to_string() for v in m.global_all] == [f"/job:ps/task:{i % 2}" for i in range(5)]
    assert all(v.device.to_string() == "/job:worker/task:1/gpu:0" for v in m.local_all)
    assert [(l.name, gl.name) for l, gl in m.sync_pairs()][0] == ("local/dense/kernel", "global/dense/kernel")
    assert m.global_trainable[0].initializer.kind == "random_normal" and m.global_trainable[0].initializer.stddev == 1.0
    assert m.global_trainable[1].initializer.kind == "zeros"
    # get_vars matches a regex prefix like tf.get_collection(key, scope)
    assert len(get_vars(m.store, "glo")) == 4 and len(get_vars(m.store, "local/dense_1")) == 2
    st = VariableStore()
    with st.variable_scope("a"):
        st.get_variable("x", [2], initializer=random_normal_initializer())
        with pytest.raises(ValueError):
            st.get_variable("x", [2])


def test_memory_fraction_formula():
    assert memory.per_process_gpu_memory_fraction(2, 1) == pytest.approx(0.45)
    assert memory.per_process_gpu_memory_fraction(16, 4) == pytest.approx(0.225)
    assert memory.per_process_gpu_memory_fraction(3, 2) == pytest.approx(0.45)
    b = memory.banner(0.45).splitlines()
    assert b[0] == "-" * 100 and b[1] == "Per-process GPU memory fraction: 0.45" and b[2] == "-" * 100


def test_launcher_task_commands_modulo_mapping():
    from distributedtensorflowexample_b200.launch import launcher
    tasks = launcher.task_commands(16, 4, ["0", "1", "2", "3"])
    assert [t["name"] for t in tasks][:3] == ["ps", "worker0", "worker1"] and len(tasks) == 17
    # the task's own GPU is GPU_ID[i % num_gpus] and comes first (= its cuda:0, like the reference's gpu:0); the
    # cluster's other GPUs follow so that shards living there can be mapped (CUDA IPC needs the device visible)
    vis = [t["env"]["CUDA_VISIBLE_DEVICES"].split(",") for t in tasks[1:]]
    assert [v[0] for v in vis] == [str(i % 4) for i in range(16)]
    assert all(sorted(v) == ["0", "1", "2", "3"] for v in vis)
    assert launcher.task_commands(2, 1, ["5"])[1]["env"]["CUDA_VISIBLE_DEVICES"] == "5"
    w5 = tasks[6]["argv"]
    assert w5[w5.index("--job_name") + 1] == "worker" and w5[w5.index("--task_index") + 1] == "5"
    assert "--task_index=0" in tasks[0]["argv"]
    cpu = launcher.task_commands(2, 1, ["0"], device="cpu")
    assert all(t["env"]["CUDA_VISIBLE_DEVICES"] == "" for t in cpu)        # like the reference's ps
    cmds = launcher.launch_tmux("sess", tasks[:3], dry_run=True)
    assert cmds[0] == ["tmux", "kill-session", "-t", "sess"]
    assert cmds[1][:6] == ["tmux", "new-session", "-s", "sess", "-n", "ps"]
    assert cmds[2][:2] == ["tmux", "new-window"] and cmds[-1][:2] == ["tmux", "a"]
    with pytest.raises(ValueError):
        launcher.task_commands(2, 4, ["0"])


def test_shell_launchers_keep_reference_knobs():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run(["bash", os.path.join(root, "run_multi_gpu.sh"), "--mode", "print"], capture_output=True,
                         text=True, env={**os.environ, "MODE": "print"}, timeout=60).stdout
    assert out.count("new-window") == 16 and "--num_workers 16 --num_gpus 4" in out
    assert "CUDA_VISIBLE_DEVICES=3,0,1,2" in out and "--task_index 15" in out
    out = subprocess.run(["bash", os.path.join(root, "run_single_gpu.sh")], capture_output=True, text=True,
                         env={**os.environ, "MODE": "print"}, timeout=60).stdout
    assert out.count("new-window") == 2 and "--num_workers 2 --num_gpus 1" in out


def test_shard_layout_offsets_pitch_and_partitioning():
    from distributedtensorflowexample_b200.models.mlp import MLPSpec, reference_mlp_spec
    from distributedtensorflowexample_b200.parallel.layout import ShardLayout, HEADER_BYTES
    from distributedtensorflowexample_b200.train.replica import build_replicated_model
    m = build_replicated_model(reference_mlp_spec(), num_ps=1)
    lay = ShardLayout(m.global_all, 1)
    offs = [lay[n].offset for n in lay.names()]
    assert offs[0] == HEADER_BYTES and all(o % 256 == 0 for o in offs) and offs == sorted(offs)
    assert lay["global/dense/kernel"].nbytes == 313600 and lay["global/global_step"].nbytes == 4
    assert lay.shard_bytes[0] >= offs[-1] + 4
    # bf16 matrices with a width that is not a multiple of 8 get a padded pitch (TMA needs 16-byte rows)
    wide = build_replicated_model(MLPSpec((784, 64, 10), param_dtype="bfloat16"), num_ps=4)
    lay = ShardLayout(wide.global_all, 4, partition_min_bytes=64 << 10)
    e = lay["global/dense_1/kernel"]
    assert e.pitch == 16 and e.nbytes == 64 * 16 * 2 and not e.parts
    w1 = lay["global/dense/kernel"]                      # 784x64 bf16 = 100352 B >= 64 KB, 784 % 4 == 0
    assert w1.shard == -1 and [p[0] for p in w1.parts] == [0, 1, 2, 3] and w1.nbytes == 784 * 64 * 2
    assert all(any(x.name == "global/dense/kernel" for x in lay.shard_entries(k)) for k in range(4))
    with pytest.raises(ValueError):
        ShardLayout(wide.local_all, 4)                   # local/* variables are not on the ps job


# ---- property tests ------------------------------------------------------------------------------------------
from hypothesis import given, settings, strategies as st


@settings(max_examples=60, deadline=None)
@given(num_workers=st.integers(1, 32), num_ps=st.integers(1, 8), base=st.integers(1024, 60000))
def test_cluster_spec_property(num_workers, num_ps, base):
    from distributedtensorflowexample_b200.cluster import cluster_spec
    spec = cluster_spec(num_workers, num_ps, base_port=base)
    hosts = spec["ps"] + spec["worker"]
    assert len(spec["ps"]) == num_ps and len(spec["worker"]) == num_workers
    ports = [int(h.rsplit(":", 1)[1]) for h in hosts]
    assert ports == list(range(base, base + num_ps + num_workers))          # consecutive, ps first
    assert all(h.startswith("127.0.0.1:") for h in hosts)


@settings(max_examples=60, deadline=None)
@given(num_workers=st.integers(1, 64), num_gpus=st.integers(1, 8))
def test_memory_fraction_property(num_workers, num_gpus):
    import math
    from distributedtensorflowexample_b200.utils.memory import per_process_gpu_memory_fraction as gpu_memory_fraction
    if num_gpus > num_workers:
        num_gpus = num_workers
    f = gpu_memory_fraction(num_workers, num_gpus)
    per_gpu = math.ceil(num_workers / num_gpus)
    assert f == 0.9 / per_gpu and f * per_gpu <= 0.9 + 1e-12                 # the workers of one GPU never exceed 90 %


@settings(max_examples=80, deadline=None)
@given(dims=st.lists(st.integers(1, 300), min_size=2, max_size=5), num_ps=st.integers(1, 6),
       dtype=st.sampled_from(["float32", "bfloat16"]), part_kb=st.sampled_from([0, 1, 16, 256]),
       strategy=st.sampled_from(["round_robin", "greedy_bytes"]))
def test_shard_layout_property(dims, num_ps, dtype, part_kb, strategy):
    """No two variables (or parts) overlap inside a shard, everything is 256-byte aligned and inside the shard,
    a partitioned variable has exactly one equally sized part on every shard."""
    from distributedtensorflowexample_b200.models.mlp import MLPSpec
    from distributedtensorflowexample_b200.parallel.layout import ShardLayout, HEADER_BYTES, ALIGN
    from distributedtensorflowexample_b200.train.replica import build_replicated_model
    m = build_replicated_model(MLPSpec(tuple(dims), param_dtype=dtype), num_ps=num_ps, ps_strategy=strategy)
    lay = ShardLayout(m.global_all, num_ps, partition_min_bytes=part_kb << 10)
    assert sorted(lay.names()) == sorted(v.name for v in m.global_all)
    spans = {k: [] for k in range(num_ps)}
    for name in lay.names():
        e = lay[name]
        if e.parts:
            assert e.shard == -1 and [pk for pk, _ in e.parts] == list(range(num_ps)) and e.nbytes % num_ps == 0
            for pk, off in e.parts:
                spans[pk].append((off, off + e.nbytes // num_ps, name))
        else:
            assert 0 <= e.shard < num_ps
            spans[e.shard].append((e.offset, e.offset + e.nbytes, name))
        item = 2 if e.dtype == "bfloat16" else 4
        rows, pitch = (e.shape[0], e.pitch or e.shape[1]) if len(e.shape) == 2 else (1, int(__import__('math').prod(e.shape)))
        assert e.nbytes == rows * pitch * item
        if e.dtype == "bfloat16" and len(e.shape) == 2:
            assert (pitch * item) % 16 == 0                                   # TMA row pitch
    for k, sp in spans.items():
        sp.sort()
        assert all(a % ALIGN == 0 and a >= HEADER_BYTES and b <= lay.shard_bytes[k] for a, b, _ in sp)
        assert all(sp[i][1] <= sp[i + 1][0] for i in range(len(sp) - 1)), sp
        assert lay.shard_bytes[k] % ALIGN == 0


@settings(max_examples=100, deadline=None)
@given(bs=st.integers(1, 10 ** 6), lr=st.floats(1e-6, 10.0, allow_nan=False), logdir=st.text(alphabet="abc/._-XYZ019", min_size=1, max_size=20),
       job=st.sampled_from(["ps", "worker"]), eq=st.lists(st.booleans(), min_size=4, max_size=4),
       noise=st.lists(st.sampled_from(["--unknown_a=1", "--unknown_b", "stray"]), max_size=3), order=st.permutations(range(4)))
def test_flag_parser_property(bs, lr, logdir, job, eq, noise, order):
    """Both spellings (`--k v`, `--k=v`), any order, unknown arguments in between: every known flag gets its
    value, every unknown argument comes back untouched and in order."""
    fv = _flags()
    known = [("batch_size", str(bs)), ("learning_rate", repr(lr)), ("logdir", logdir), ("job_name", job)]
    argv = []
    for i in order:
        k, v = known[i]
        argv += [f"--{k}={v}"] if eq[i] else [f"--{k}", v]
        if noise and i % 2 == 0:
            argv.append(noise[i % len(noise)])
    rest = fv.parse(list(argv))
    assert fv.batch_size == bs and fv.learning_rate == lr and fv.logdir == logdir and fv.job_name == job
    assert rest == [a for a in argv if a in ("--unknown_a=1", "--unknown_b", "stray")]


def test_help_prints_flags_and_exits_zero(capsys):
    fv = _flags()
    with pytest.raises(SystemExit) as e:
        fv.parse(["--job_name", "worker", "--help"])
    assert e.value.code == 0
    out = capsys.readouterr().out
    assert "--learning_rate (float, default 0.001)" in out and "--job_name (string, default 'ps')" in out
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "main.py", "--help"], cwd=root, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and "--training_steps" in r.stdout        # and does not start the default ps role


def test_shutdown_signal_ends_a_task_stuck_in_native_code(tmp_path):
    """SIGTERM while the main thread cannot run Python handlers (signals masked, like during a gloo exchange, or
    a native rendezvous that never returns): the watchdog exits with the handler's code after the grace period."""
    import signal as sig
    import time
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = ("import signal, time, sys\n"
            f"sys.path.insert(0, {root!r})\n"
            "from distributedtensorflowexample_b200.utils.signals import install_shutdown_handlers, HANDLED\n"
            "install_shutdown_handlers(grace_s=1.0)\n"
            "signal.pthread_sigmask(signal.SIG_BLOCK, set(HANDLED))\n"
            "print('ready', flush=True)\n"
            "time.sleep(120)\n")
    p = subprocess.Popen([sys.executable, "-c", code], stdout=subprocess.PIPE, text=True)
    assert p.stdout.readline().strip() == "ready"
    t0 = time.time()
    p.send_signal(sig.SIGTERM)
    assert p.wait(timeout=30) == 143 and time.time() - t0 < 15
    # and the ordinary path is unchanged: an interruptible main thread leaves through SystemExit at once
    code2 = code.replace("signal.pthread_sigmask(signal.SIG_BLOCK, set(HANDLED))\n", "")
    p = subprocess.Popen([sys.executable, "-c", code2], stdout=subprocess.PIPE, text=True)
    assert p.stdout.readline().strip() == "ready"
    t0 = time.time()
    p.send_signal(sig.SIGINT)
    assert p.wait(timeout=30) == 130 and time.time() - t0 < 0.9


def test_launcher_cli_forwards_flags_after_a_double_dash(capsys):
    from distributedtensorflowexample_b200.launch import launcher
    assert launcher.main(["--mode", "print", "--num_workers", "2", "--device", "cpu", "--", "--training_steps", "7"]) == 0
    out = capsys.readouterr().out
    assert "--training_steps 7" in out and " -- " not in out         # a bare '--' would end flag parsing in every task
