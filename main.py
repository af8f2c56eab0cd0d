#!/usr/bin/env python
"""Entry point with the reference's role switch and flags (reference main.py:26-80).

    python main.py --job_name ps     --task_index 0 --num_workers 2 --num_gpus 1
    python main.py --job_name worker --task_index 0 --num_workers 2 --num_gpus 1

Every process loads the data set, builds the cluster spec, installs the signal handlers and
prints the GPU memory fraction banner before branching on its role - the ps included, as in the
reference.  Use run_single_gpu.sh / run_multi_gpu.sh to start a whole cluster.
"""
from __future__ import annotations

import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from distributedtensorflowexample_b200 import flags as flags_lib
from distributedtensorflowexample_b200.cluster import ClusterSpec, cluster_spec
from distributedtensorflowexample_b200.utils import memory, signals

flags_lib.define_reference_flags()
flags_lib.define_framework_flags()
FLAGS = flags_lib.FLAGS


def main() -> int:
    import faulthandler, signal as _signal
    faulthandler.register(_signal.SIGUSR1, all_threads=True)     # kill -USR1 <pid> dumps every thread's stack
    from distributedtensorflowexample_b200.data import find_data_dir, read_data_sets
    from distributedtensorflowexample_b200.train.server import Server, ServerConfig, resolve_backend

    # Load MNIST dataset (one_hot=True like the reference; ids are derived where the GPU wants them).
    seed = None if FLAGS.seed < 0 else FLAGS.seed + 17 * FLAGS.task_index + (0 if FLAGS.job_name == "ps" else 1)
    mnist = read_data_sets(find_data_dir(FLAGS.data_dir), one_hot=True, seed=seed,
                           synthetic=(FLAGS.data == "synthetic"))

    # Cluster specification
    cluster = ClusterSpec(cluster_spec(FLAGS.num_workers, FLAGS.num_ps, base_port=FLAGS.base_port))

    # Signals: exit code 128+signum
    signals.install_shutdown_handlers()

    # GPU memory fraction (computed and printed by every role)
    fraction = memory.per_process_gpu_memory_fraction(FLAGS.num_workers, FLAGS.num_gpus)
    print(memory.banner(fraction), flush=True)

    backend = resolve_backend(FLAGS.device)
    if FLAGS.job_name == "ps":
        server = Server(cluster, job_name="ps", task_index=FLAGS.task_index,
                        config=ServerConfig(backend=backend))
        print(f"ps:{FLAGS.task_index} serving at {server.target} (data source: {mnist.source})", flush=True)
        die_after = None                  # test-only fault injection: --fault kill_ps:<seconds>
        for part in (FLAGS.fault or "").split(","):
            if part.startswith("kill_ps:"):
                die_after = time.time() + float(part.split(":", 1)[1])
        while True:                       # not server.join(): signals must stay deliverable
            time.sleep(0.5)
            if die_after is not None and time.time() >= die_after:
                print("fault injection: ps dies", flush=True)
                os._exit(19)
            if server.error is not None:
                raise server.error
            if backend == "gloo" and not server.serving():
                return 0                  # every worker said goodbye (headless launcher / tests)
    elif FLAGS.job_name == "worker":
        from distributedtensorflowexample_b200.train.worker import Worker
        server = Server(cluster, job_name="worker", task_index=FLAGS.task_index,
                        config=ServerConfig(backend=backend, gpu_memory_fraction=fraction))
        worker = Worker(FLAGS.job_name, FLAGS.task_index, server)
        try:
            worker.learn(mnist)
        finally:
            if backend == "gloo":
                try:
                    worker.client.shutdown_servers()
                except Exception:
                    pass
        return 0
    raise SystemExit(f"--job_name must be 'ps' or 'worker', got {FLAGS.job_name!r}")


def _run() -> None:
    code = 0
    try:
        code = main() or 0
    except SystemExit as e:          # 128+signum from the handlers, or an argument error
        code = e.code if isinstance(e.code, int) else (0 if e.code is None else 1)
        if not isinstance(e.code, int) and e.code is not None:
            print(e.code, file=sys.stderr)
    finally:
        sys.stdout.flush()
        sys.stderr.flush()
    # Leave without running the collective back-end's destructors: with a peer already gone they can
    # wait for their full timeout, and the exit code (e.g. 143 after SIGTERM) is what callers observe.
    os._exit(code)


if __name__ == "__main__":
    _run()
