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
On the GPU path ``steps_per_launch > 1`` fuses several steps into one kernel launch; the cadence
rules are then evaluated for every step of the launch after it returns.
"""
from __future__ import annotations

import os
import time

import numpy as np

from ..flags import FLAGS
from ..models.mlp import spec_from_flags
from ..utils.checkpoint import FastSaver
from ..utils.events import FileWriter, training_summary
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

                def exchange(blob):
                    import pickle
                    kv.set(f"sync/{me}", pickle.dumps(blob))
                    return [pickle.loads(kv.get(f"sync/{r}")) for r in range(W)]
                # variable on ps shard k is reduced by worker (k mod W)
                # a round may have to wait for workers that are still polling for the chief's initialisation
                self.sync_group = SyncGroup(store, names, me, W, exchange=exchange,
                                            timeout_s=f.sync_timeout_secs + 2.0 * f.recovery_wait_secs)
                kw["grad_buffers"] = self.sync_group.grad_views({n: tuple(store.tensor(n).shape) for n in names})
            else:
                kw["adam"] = dict(slots=[store.tensor(n + "/grad_slot") for n in names],
                                  m=[store.tensor(n + "/Adam") for n in names],
                                  v=[store.tensor(n + "/Adam_1") for n in names], apply=self.is_chief)
            fused_ok = (f.sync and f.optimizer == "sgd" and f.engine in ("auto", "fused")
                        and FusedMlp2Stepper.supports(self.spec, f.batch_size))
            tr = GpuTrainer(self.spec, f.batch_size, f.learning_rate, [store.tensor(n) for n in names],
                            store.tensor(self.global_step.name), dev, engine="fused" if fused_ok else "layerwise",
                            steps_per_launch=f.steps_per_launch if fused_ok else 1, remote=True, tf32=f.tf32)
            if fused_ok:
                # the whole synchronous step - exchange included - is one kernel (mlp2_step.cu phase D)
                tr.fused = FusedMlp2Stepper(self.spec, f.batch_size, tr.params, tr.global_step, f.learning_rate, dev,
                                            max_steps_per_launch=max(1, f.steps_per_launch), remote=True, **kw)
                tr.fused.sync_exchange = self.sync_group.exchange_struct(f.learning_rate)
                self._sync_in_kernel = True
            else:
                tr.layer = LayerwiseStepper(self.spec, f.batch_size, tr.params, tr.global_step, f.learning_rate, dev,
                                            atomic=not f.unlocked_apply, tf32=f.tf32, **kw)
            self.trainer = tr
        else:
            self.trainer = GpuTrainer(self.spec, f.batch_size, f.learning_rate,
                                      [store.tensor(n) for n in names], store.tensor(self.global_step.name), dev,
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

    def _after_step(self, step: int, cost: float, acc: float, eval_fn, write_summary: bool = True) -> bool:
        """Summary + cadence rules of reference worker.py:139-159. Returns True to stop."""
        f = self.flags
        if write_summary:
            with self.tracer.span("summary", "host"):
                self.summary_writer.add_summary(training_summary(cost, acc), step)
        if "kill_worker" in self.faults and self.faults["kill_worker"][0] == self.task_index \
                and step >= self.faults["kill_worker"][1]:
            print(f"fault injection: worker {self.task_index} dies at step {step}", flush=True)
            self.summary_writer.flush()
            os._exit(17)
        if step % f.log_every == 0 and step != 0:
            elapsed = time.time() - self._start_time
            print("step: {}\t| cost: {}\t| speed: {}step/sec".format(
                step, cost, float((step - self._start_step) / max(elapsed, 1e-9))), flush=True)
            self._start_time = time.time()
            self._start_step = step
        if step % f.eval_every == 0:
            with self.tracer.span("eval", "host", step=step):
                print("test accuracy: {}".format(eval_fn()), flush=True)
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
        train = DeviceDataset(dataset.train.images, dataset.train.label_ids, dev, seed=seed)
        test_x = torch.as_tensor(dataset.test.images, dtype=torch.float32).to(dev)
        test_y = torch.as_tensor(dataset.test.label_ids, dtype=torch.int32).to(dev)
        eval_fn = lambda: tr.evaluate(test_x, test_y)
        separate_exchange = self.sync_group is not None and not self._sync_in_kernel
        n = 1 if separate_exchange else max(1, f.steps_per_launch)
        step = 0
        while not sv.should_stop():
            with tc.span("launch", "step", steps=n), tc.device_span("step_kernel", steps=n):
                loss_d, acc_d, step_d = tr.train_device_steps(train, n)
            if separate_exchange:
                with tc.span("sync_exchange", "step"):
                    self.sync_group.reduce_scatter_apply(f.learning_rate)
                step_d = self.store.tensor(self.global_step.name)[:1] - 1      # the aggregated update just applied
            if self.sync_group is not None and self.sync_group.aborted():
                print("sync: a peer left, stopping", flush=True)
                break
            with tc.span("results_d2h", "step"):
                if tr.fused is not None and not separate_exchange:
                    loss, acc, steps = tr.fused.k.results_host(len(loss_d))        # one copy, one sync per launch
                else:
                    loss, acc, steps = loss_d.tolist(), acc_d.tolist(), step_d.tolist()
            # per-step summaries of the whole launch in one native call, up to the step that ends training
            m = next((i + 1 for i, s in enumerate(steps) if s >= f.training_steps), len(steps))
            with tc.span("summary", "host", steps=m):
                self.summary_writer.add_training_summaries(steps[:m], loss[:m], acc[:m])
            stop = False
            for l, a, s in zip(loss[:m], acc[:m], steps[:m]):
                step = s
                if self._after_step(s, l, a, eval_fn, write_summary=False):
                    stop = True
                    break
            if stop:
                break
        return step
