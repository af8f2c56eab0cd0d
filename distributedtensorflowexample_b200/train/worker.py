"""The worker: replicated-model construction and the training loop (reference worker.py:7-159;
SURVEY C8-C17, 3.3, 3.4).

``Worker(job_name, task_index, server)`` mirrors the reference constructor; ``learn(dataset)`` is the
same loop: pull -> batch -> forward/backward/push -> summary -> counter -> log/eval cadence ->
``training_steps`` bound, under a Supervisor-managed session.

Observable behaviour kept from the reference:
  * ``step`` is the GLOBAL step observed during the train op, pre-increment (worker.py:136);
  * ``step: {}\\t| cost: {}\\t| speed: {}step/sec`` when ``step % 100 == 0 and step != 0`` (speed =
    global steps/sec as seen by this worker, first interval from session start, worker.py:143-148);
  * ``test accuracy: {}`` on the full test set when ``step % 10000 == 0`` - including step 0
    (worker.py:150-154);
  * a loss/accuracy summary per step, keyed by the observed global step, in ``<logdir>_<task>``
    (worker.py:96,139);
  * stop when ``step >= training_steps`` (a bound on the global step, worker.py:158-159).
On the GPU path several steps are fused into one kernel launch (``--steps_per_launch``, adaptive by default);
the cadence rules are evaluated for every step of a launch when it returns, launch sizes are planned so that
evaluation and the end of training fall on launch boundaries (``plan_launch_steps``), two launches are kept in
flight, and the per-step summaries are written by a helper thread.
"""
from __future__ import annotations

import os
import time

import numpy as np

from ..flags import FLAGS
from ..models.mlp import spec_from_flags
from ..utils.checkpoint import FastSaver
from ..utils.events import FileWriter, Summary, TAG_TEST_ACCURACY, training_summary
from ..utils import tracing
from ..parallel.gloo_ps import SyncAborted
from .replica import build_replicated_model
from .supervisor import Supervisor


def _parse_fault(text: str):
    """'kill_worker:1@50,delay_push:0.01' -> {'kill_worker': (1, 50), 'delay_push': 0.01}"""
    out = {}
    for part in (text or "").split(","):
        part = part.strip()
        if not part:
            continue
        name, _, arg = part.partition(":")
        if name == "kill_worker":
            task, _, step = arg.partition("@")
            out[name] = (int(task), int(step))
        elif name == "delay_push":
            out[name] = float(arg)
        elif name == "kill_ps":
            out[name] = float(arg)          # handled by the ps process (main.py)
        else:
            raise ValueError(f"unknown fault {part!r}")
    return out


AUTO_STEPS_PER_LAUNCH = 256     # --steps_per_launch 0: cap of the adaptive launch size
PIPELINE_DEPTH = 2              # launch groups in flight: the GPU runs one while the host digests the previous one


def plan_launch_steps(next_step_lb: int, training_steps: int, eval_every: int, num_workers: int, cap: int,
                      nothing_in_flight: bool) -> int:
    """How many steps the next kernel launch of a worker should fuse (0: enqueue nothing, wait for results).

    ``next_step_lb`` is the estimate of the global step the first step of the launch will observe: last observed
    step + 1 + ``num_workers`` x this worker's steps in flight (exact for a lone worker; with peers, every step of
    ours that is still queued stands for about ``num_workers`` global steps).  The reference checks its
    cadence rules after every single step (worker.py:143-159); a multi-step launch checks them when it returns,
    so its size is chosen such that, counting on this worker's own steps alone,
      * the step that observes ``training_steps`` (the last one the reference would run) is the last of a launch,
      * so is a step that observes a multiple of ``eval_every`` (the evaluation then sees the parameters right
        after that step's update),
    and with ``num_workers`` workers racing - the global step then advances about ``num_workers`` times faster than
    this worker's own count - launches shrink geometrically towards those boundaries instead of running over them.
    """
    W = max(1, num_workers)
    rem = training_steps - next_step_lb + 1            # own steps until one observes >= training_steps, inclusive
    if rem <= 0:
        # every step enqueued from now on would observe >= training_steps: only the reference's "one more step,
        # then notice" remains, and only if no step in flight is going to notice first
        return 1 if nothing_in_flight else 0
    n = min(cap, -(-rem // W))
    if eval_every > 0:
        boundary = -(-next_step_lb // eval_every) * eval_every
        n = min(n, -(-(boundary - next_step_lb + 1) // W))
    return max(1, n)


class Worker:
    def __init__(self, job_name: str, task_index: int, server, flags=FLAGS):
        self.job_name, self.task_index, self.server, self.flags = job_name, task_index, server, flags
        self.is_chief = task_index == 0
        self.faults = _parse_fault(flags.fault)
        seed = None if flags.seed < 0 else flags.seed
        self.rng = np.random.default_rng(seed)
        self.spec = spec_from_flags(flags.hidden, dtype=flags.dtype if server.backend == "cuda" else "float32")
        # C8: two copies of the net, global on the ps job, local on this worker
        self.model = build_replicated_model(self.spec, job_name, task_index, num_ps=server.num_ps,
                                            ps_strategy=flags.ps_strategy, optimizer=flags.optimizer)
        self.global_step = self.model.global_step
        if flags.log_device_placement:
            for v in self.model.store.all_variables():
                print(f"{v.name}: ({'VariableV2'}) {v.device}", flush=True)
        if server.backend == "gloo":
            self._build_cpu()
        else:
            self._build_cuda()
        self.build_summary_op()
        self.build_saver()

    # ---- construction -------------------------------------------------------------------------
    def _build_cpu(self) -> None:
        from ..parallel.gloo_ps import GlooParamClient
        from .cpu_step import CpuStepEngine
        self.client = GlooParamClient(self.model.global_all, self.server.cluster)
        self.engine = CpuStepEngine(self.model, self.client, self.flags.learning_rate,
                                    push_delay_s=self.faults.get("delay_push", 0.0),
                                    optimizer=self.flags.optimizer, sync=self.flags.sync,
                                    sync_timeout_s=self.flags.sync_timeout_secs,
                                    first_round_grace_s=2.0 * self.flags.recovery_wait_secs)
        self.trainer = None

    def _build_cuda(self) -> None:
        import torch
        from ..parallel.client import PeerParamClient
        from ..parallel.ps import PeerParamStore
        from .trainer import GpuTrainer
        dev = torch.device("cuda", self.server.config.device_index)
        store = PeerParamStore(self.model.global_all, self.server.num_ps, dev)
        kv = self.server.store
        if self.is_chief:   # the chief tells each ps how much state to host
            for k in range(self.server.num_ps):
                kv.set(f"shard/{k}/bytes", str(store.layout.shard_bytes[k]))
        handles = {k: kv.get(f"shard/{k}/handle") for k in range(self.server.num_ps)}   # blocks until exported
        store.open_handles(handles)
        self.store = store
        self.client = PeerParamClient(store)
        names = [v.name for v in self.model.global_trainable]
        f = self.flags
        self.sync_group = None
        self._sync_in_kernel = False
        if f.sync or f.optimizer == "adam":
            from .gpu_step import FusedMlp2Stepper, LayerwiseStepper
            kw = {}
            if f.sync:
                if self.server.num_workers > max(1, f.num_gpus):
                    raise ValueError("--sync needs one worker per GPU (its kernel waits on the other ranks)")
                from ..parallel.sync import SyncGroup
                W, me = self.server.num_workers, self.task_index

                # Workers may be restarted against a ps that stayed up: key the handle exchange by generation
                # (every W joins start a new one) so that nobody opens a dead process's IPC handles, and let the
                # generation's chief clear the stop word its predecessors raised on their way out.
                gen = (kv.add("sync/joined", 1) - 1) // W

                def exchange(blob):
                    import pickle
                    if me == 0:
                        store.clear_stop()
                    kv.set(f"sync/{gen}/{me}", pickle.dumps(blob))
                    return [pickle.loads(kv.get(f"sync/{gen}/{r}")) for r in range(W)]
                # variable on ps shard k is reduced by worker (k mod W)
                # a round may have to wait for workers that are still polling for the chief's initialisation
                fused_sync = (f.optimizer == "sgd" and f.engine in ("auto", "fused")
                              and FusedMlp2Stepper.supports(self.spec, f.batch_size))
                self.sync_group = SyncGroup(store, names, me, W, exchange=exchange,
                                            timeout_s=f.sync_timeout_secs + 2.0 * f.recovery_wait_secs,
                                            mode=f.sync_mode if fused_sync else "reduce_scatter")
                kw["grad_buffers"] = self.sync_group.grad_views({n: tuple(store.tensor(n).shape) for n in names})
            else:
                kw["adam"] = dict(slots=[store.tensor(n + "/grad_slot") for n in names],
                                  m=[store.tensor(n + "/Adam") for n in names],
                                  v=[store.tensor(n + "/Adam_1") for n in names],
                                  t=store.tensor("global/adam_t"), apply=self.is_chief)
            fused_ok = (f.sync and f.optimizer == "sgd" and f.engine in ("auto", "fused")
                        and FusedMlp2Stepper.supports(self.spec, f.batch_size))
            # fused synchronous mode trains on this rank's own copies (the shard where it owns the variable)
            params = self.sync_group.home_params() if fused_ok else [store.tensor(n) for n in names]
            gstep = self.sync_group.home(self.global_step.name) if fused_ok else store.tensor(self.global_step.name)
            tr = GpuTrainer(self.spec, f.batch_size, f.learning_rate, params, gstep, dev,
                            engine="fused" if fused_ok else "layerwise",
                            steps_per_launch=f.steps_per_launch if fused_ok else 1, remote=True, tf32=f.tf32)
            if fused_ok:
                # the whole synchronous step - exchange included - is one kernel (mlp2_step.cu phase D)
                tr.fused = FusedMlp2Stepper(self.spec, f.batch_size, tr.params, tr.global_step, f.learning_rate, dev,
                                            max_steps_per_launch=max(1, f.steps_per_launch), remote=False, **kw)
                tr.fused.sync_exchange = self.sync_group.exchange_struct(f.learning_rate)
                self._sync_in_kernel = True
            else:
                tr.layer = LayerwiseStepper(self.spec, f.batch_size, tr.params, tr.global_step, f.learning_rate, dev,
                                            atomic=not f.unlocked_apply, tf32=f.tf32, **kw)
            self.trainer = tr
        else:
            self.trainer = GpuTrainer(self.spec, f.batch_size, f.learning_rate,
                                      [store.param_handle(n) for n in names], store.tensor(self.global_step.name), dev,
                                      steps_per_launch=f.steps_per_launch, engine=f.engine,
                                      atomic=not f.unlocked_apply, remote=True, tf32=f.tf32)
        self.engine = None

    def build_summary_op(self) -> None:
        """reference worker.py:87-96: the writer (and its event file) exists from construction on."""
        self.summary_writer = FileWriter(self.flags.logdir + "_%d" % self.task_index)

    def build_saver(self) -> None:
        """reference worker.py:102-103: FastSaver over the 4 global weights + global_step."""
        self.saver = FastSaver([v.name for v in self.model.global_all])

    def global_init_op(self) -> None:
        self.client.initialize(self.rng)

    def local_init_op(self) -> None:
        if self.engine is not None:
            self.engine.local_init_op(self.rng)

    def graph_text(self) -> str:
        lines = [f"# between-graph replica of worker {self.task_index}; model dims {list(self.spec.dims)}"]
        for v in self.model.store.all_variables():
            lines.append(f'node {{ name: "{v.name}" op: "VariableV2" device: "{v.device}" '
                         f'attr {{ shape: {list(v.shape)} dtype: "{v.dtype}" trainable: {str(v.trainable).lower()} }} }}')
        return "\n".join(lines) + "\n"

    # ---- the loop -------------------------------------------------------------------------------
    def learn(self, dataset) -> int:
        f = self.flags
        sv = Supervisor(is_chief=self.is_chief, logdir=f.logdir, saver=self.saver, client=self.client,
                        summary_writer=self.summary_writer, save_model_secs=f.save_model_secs,
                        save_summaries_secs=f.save_summaries_secs, init_op=self.global_init_op,
                        local_init_op=self.local_init_op, recovery_wait_secs=f.recovery_wait_secs,
                        graph_text=self.graph_text())
        self.sv = sv
        last_step = 0
        if f.trace_file:
            tracing.enable(process_name=f"worker:{self.task_index}")
        self.tracer = tracing.get_tracer()
        with self.tracer.span("managed_session", "session"), sv.managed_session() as sess:
            if sv.restored_from:
                print(f"restored from {sv.restored_from}", flush=True)
            if getattr(self, "_sync_in_kernel", False):
                self.sync_group.refresh_replicas()      # the model is ready: load this rank's copies from the shards
            self._begin_time = time.time()
            self._start_time = time.time()
            self._start_step = 0
            run = self._loop_cpu if self.trainer is None else self._loop_cuda
            try:
                last_step = run(sv, dataset)
            finally:
                if getattr(self, "sync_group", None) is not None or (self.trainer is None and f.sync):
                    # tell the peers' spinning sync kernels (CUDA) / the ps holding their round (gloo) that this
                    # rank is gone, so nobody waits for a round that cannot complete
                    try:
                        self.client.request_stop()
                    except Exception:
                        pass
        self.summary_writer.close()
        if f.trace_file:
            path = self.tracer.save(tracing.trace_path_for_task(f.trace_file, "worker", self.task_index))
            print(f"trace written to {path}", flush=True)
        return last_step

    def _after_step(self, step: int, cost: float, acc: float, eval_fn, write_summary: bool = True,
                    now: float = None) -> bool:
        """Summary + cadence rules of reference worker.py:139-159. Returns True to stop.  ``now``: when the step
        finished (steps of a multi-step launch are stamped by interpolation over the launch)."""
        f = self.flags
        if write_summary:
            with self.tracer.span("summary", "host"):
                self.summary_writer.add_summary(training_summary(cost, acc), step)
        if "kill_worker" in self.faults and self.faults["kill_worker"][0] == self.task_index \
                and step >= self.faults["kill_worker"][1]:
            print(f"fault injection: worker {self.task_index} dies at step {step}", flush=True)
            if getattr(self, "_summary_thread", None) is not None:
                self._summary_thread.drain()
            self.summary_writer.flush()
            os._exit(17)
        if f.log_every > 0 and step % f.log_every == 0 and step != 0:
            now = time.time() if now is None else now
            elapsed = now - self._start_time
            print("step: {}\t| cost: {}\t| speed: {}step/sec".format(
                step, cost, float((step - self._start_step) / max(elapsed, 1e-9))), flush=True)
            self._start_time = now
            self._start_step = step
        if f.eval_every > 0 and step % f.eval_every == 0:
            with self.tracer.span("eval", "host", step=step):
                test_acc = eval_fn()
                print("test accuracy: {}".format(test_acc), flush=True)
            # additive: the evaluation also goes to the event file (the reference only prints it), so a run's
            # time-to-accuracy can be read from its logs: wall_time of this event minus that of the first step's
            if getattr(self, "_summary_thread", None) is not None:
                self._summary_thread.drain()         # keep the file in step order
            self.summary_writer.add_summary(Summary([(TAG_TEST_ACCURACY, float(test_acc))]), step)
        return step >= f.training_steps

    def _loop_cpu(self, sv: Supervisor, dataset) -> int:
        eng, f = self.engine, self.flags
        step = 0
        while not sv.should_stop():
            tr = self.tracer
            with tr.span("sync_op(pull)", "step"):
                eng.sync_op()                                                    # reference :131
            batch_x, batch_y = dataset.train.next_batch(f.batch_size)           # :133
            try:
                with tr.span("train_op(fwd/bwd/push)", "step"):
                    cost, acc, step = eng.train_op(batch_x, batch_y)             # :135-137
            except SyncAborted as e:
                print(f"sync: a peer left, stopping ({e})", flush=True)
                break
            with tr.span("counter_op", "step"):
                eng.counter_op()                                                 # :141 (summary :139 is in _after_step)
            if self._after_step(step, cost, acc, lambda: eng.accuracy(dataset.test.images, dataset.test.labels)):
                break
        return step

    def _loop_cuda(self, sv: Supervisor, dataset) -> int:
        import torch
        from .gpu_step import DeviceDataset
        f, tr, tc = self.flags, self.trainer, self.tracer
        dev = tr.device
        seed = None if f.seed < 0 else f.seed + 1000 * self.task_index
        train = DeviceDataset(dataset.train.images, dataset.train.label_ids, dev, seed=seed, dtype=tr.input_dtype)
        test_x = torch.as_tensor(dataset.test.images, dtype=torch.float32).to(dev).to(tr.input_dtype)
        test_y = torch.as_tensor(dataset.test.label_ids, dtype=torch.int32).to(dev)
        eval_fn = lambda: tr.evaluate(test_x, test_y)
        if self.sync_group is not None and not self._sync_in_kernel:
            return self._loop_cuda_separate_exchange(sv, train, eval_fn)
        # One launch = up to `cap` steps of the persistent kernel (fused engine) or of the per-layer kernels; two
        # launches are kept in flight so the GPU never waits for this loop.  --steps_per_launch N pins the cap
        # (1 = the reference's one sess.run per step, at ~10x the cost per step); 0 adapts it.
        cap = f.steps_per_launch if f.steps_per_launch > 0 else AUTO_STEPS_PER_LAUNCH
        if tr.fused is None:
            cap = min(cap, 16)          # per-layer engines launch every step's kernels from this thread anyway
        W = 1 if self.sync_group is not None else self.server.num_workers   # sync: every worker observes every step
        from collections import deque
        pending = deque()
        in_flight = 0
        next_lb = self.client.read_step()       # restored runs continue from the checkpoint's step
        writer = _SummaryThread(self.summary_writer)
        self._summary_thread = writer
        step, stop, hold = 0, False, False
        t_prev = time.time()
        try:
            while not stop and not sv.should_stop():
                while len(pending) < PIPELINE_DEPTH and not hold:
                    # where the global step will be when this launch starts: every step of ours still in flight
                    # stands for about W global steps (the other workers run at our pace)
                    lb = next_lb + W * in_flight
                    n = plan_launch_steps(lb, f.training_steps, f.eval_every, W, cap, not pending)
                    if n == 0:
                        break
                    # a launch that reaches an evaluation step is digested before anything else is queued, so the
                    # evaluation reads the parameters right after that step, not a launch later
                    hold = f.eval_every > 0 and lb + n - 1 >= -(-lb // f.eval_every) * f.eval_every
                    with tc.span("launch", "step", steps=n), tc.device_span("step_kernel", steps=n):
                        pending.append(tr.launch_async(train, n, ring_depth=PIPELINE_DEPTH + 1,
                                                       ring_cap=max(cap, AUTO_STEPS_PER_LAUNCH)))
                    in_flight += n
                tk = pending.popleft()
                with tc.span("results_d2h", "step"):
                    loss, acc, steps = tr.collect(tk)
                t_now = time.time()
                in_flight -= tk.nsteps
                if self.sync_group is not None and self.sync_group.aborted():
                    print("sync: a peer left, stopping", flush=True)
                    break
                step, stop = self._digest_launch(loss, acc, steps, t_prev, t_now, eval_fn, writer)
                next_lb = max(next_lb, step + 1)
                t_prev = t_now
                if not pending:
                    hold = False
        finally:
            # launches already queued run to completion on the GPU (they cannot be recalled); their steps were
            # applied and counted, but like a reference worker that was told to stop, nobody logs them.  Wait for
            # them: the shards they push into must stay mapped until they are done.
            try:
                torch.cuda.synchronize(dev)
            except Exception:
                pass
            writer.close()
        return step

    def _digest_launch(self, loss, acc, steps, t0: float, t1: float, eval_fn, writer):
        """Per-step summaries and cadence rules (reference worker.py:139-159) for all steps of one launch, touching
        only the steps that print, evaluate or end training.  Returns (last step handled, stop)."""
        f = self.flags
        n = len(steps)
        over = np.nonzero(steps >= f.training_steps)[0]
        m = int(over[0]) + 1 if len(over) else n          # the step that ends training is the last one handled
        with self.tracer.span("summary", "host", steps=m):
            writer.submit(steps[:m], loss[:m], acc[:m])
        hot = np.zeros(m, dtype=bool)
        if f.log_every > 0:
            hot |= steps[:m] % f.log_every == 0
        if f.eval_every > 0:
            hot |= steps[:m] % f.eval_every == 0
        hot[m - 1] = True
        if "kill_worker" in self.faults and self.faults["kill_worker"][0] == self.task_index:
            hot |= steps[:m] >= self.faults["kill_worker"][1]
        stop = False
        for i in np.nonzero(hot)[0]:
            now = t0 + (t1 - t0) * (i + 1) / n
            if self._after_step(int(steps[i]), float(loss[i]), float(acc[i]), eval_fn, write_summary=False, now=now):
                stop = True
                break
        return int(steps[m - 1]), stop

    def _loop_cuda_separate_exchange(self, sv: Supervisor, train, eval_fn) -> int:
        """Synchronous mode of the per-layer engines: one step, then the stand-alone exchange kernel."""
        f, tr, tc = self.flags, self.trainer, self.tracer
        step = 0
        while not sv.should_stop():
            with tc.span("launch", "step", steps=1), tc.device_span("step_kernel", steps=1):
                loss_d, acc_d, _ = tr.train_device_steps(train, 1)
            with tc.span("sync_exchange", "step"):
                self.sync_group.reduce_scatter_apply(f.learning_rate)
            step_d = self.store.tensor(self.global_step.name)[:1] - 1      # the aggregated update just applied
            if self.sync_group.aborted():
                print("sync: a peer left, stopping", flush=True)
                break
            with tc.span("results_d2h", "step"):
                l, a, step = loss_d.item(), acc_d.item(), int(step_d.item())
            if self._after_step(step, l, a, eval_fn):
                break
        return step


class _SummaryThread:
    """Per-step summaries are written by a helper thread (the native writer releases the GIL): the training
    thread only hands over copies of a launch's result arrays, in order."""

    def __init__(self, file_writer):
        import queue
        import threading
        self._w = file_writer
        self._q = queue.Queue()
        self._err = None
        self._t = threading.Thread(target=self._run, name="summary-writer", daemon=True)
        self._t.start()

    def _run(self) -> None:
        while True:
            item = self._q.get()
            if item is None:
                return
            try:
                if callable(item):
                    item()
                else:
                    self._w.add_training_arrays(*item)
            except BaseException as e:      # surfaced by the next submit/close on the training thread
                self._err = e

    def submit(self, steps, loss, acc) -> None:
        if self._err is not None:
            raise self._err
        self._q.put((np.ascontiguousarray(steps).copy(), np.ascontiguousarray(loss).copy(),
                     np.ascontiguousarray(acc).copy()))

    def drain(self) -> None:
        import threading
        done = threading.Event()
        self._q.put(done.set)
        done.wait(30)

    def close(self) -> None:
        if self._t.is_alive():
            self._q.put(None)
            self._t.join(30)
        if self._err is not None:
            raise self._err
