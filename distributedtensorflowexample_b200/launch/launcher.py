"""Cluster launcher (reference run_single_gpu.sh / run_multi_gpu.sh).

Keeps the reference's knobs - ``name`` (tmux session), ``num_workers``, ``num_gpus``, ``GPU_ID`` array -
and its process model: one ps process per ps task + one worker process per worker task on localhost,
worker ``i`` pinned to ``GPU_ID[i % num_gpus]`` through ``CUDA_VISIBLE_DEVICES``: that GPU comes FIRST in
the task's list, so ``gpu:0`` is the task's own GPU exactly as in the reference, and the cluster's other
GPUs follow it.  They are listed only so that the task can map the parameter shards that live there:
CUDA IPC (``cudaIpcOpenMemHandle``) cannot open memory of a GPU the opening process does not see.

Two modes:
  * ``--mode tmux``     : one tmux session, one window per task, like the reference (attach with
                          ``tmux a``; ``ctrl-b <n>`` flips windows).  Teardown is manual.
  * ``--mode headless`` : the launcher is the parent of all tasks, streams their output with a
                          ``[ps0]`` / ``[worker3]`` prefix, forwards SIGINT/SIGTERM, waits for the workers,
                          then stops the ps tasks and returns the worst exit code (the supervision the
                          reference lacks).

The reference's ps runs with ``CUDA_VISIBLE_DEVICES=`` (CPU only).  Here a ps task owns an HBM
shard, so on a GPU box ps ``k`` is pinned to ``GPU_ID[k % num_gpus]``; with ``--device cpu`` (gloo
plumbing) every task runs with an empty ``CUDA_VISIBLE_DEVICES`` exactly like the reference's ps.
"""
from __future__ import annotations

import argparse
import os
import shlex
import shutil
import signal
import subprocess
import sys
import threading
import time
from typing import Dict, List, Optional, Sequence

REPO = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
MAIN = os.path.join(REPO, "main.py")


def task_commands(num_workers: int, num_gpus: int, gpu_ids: Sequence[str], num_ps: int = 1, device: str = "auto",
                  extra: Sequence[str] = (), python: str = sys.executable) -> List[Dict]:
    """One dict per task: name, env (CUDA_VISIBLE_DEVICES), argv."""
    if num_gpus < 1 or len(gpu_ids) < num_gpus:
        raise ValueError(f"GPU_ID has {len(gpu_ids)} entries but num_gpus={num_gpus}")
    common = ["--num_workers", str(num_workers), "--num_gpus", str(num_gpus), "--num_ps", str(num_ps),
              "--device", device, *extra]
    cpu_only = device == "cpu"
    tasks = []

    def visible(j: int) -> str:
        """Own GPU first (= cuda:0 of the task), then the other GPUs of the cluster for peer mapping."""
        own = str(gpu_ids[j % num_gpus])
        return ",".join([own] + [str(g) for g in gpu_ids[:num_gpus] if str(g) != own])
    for k in range(num_ps):
        vis = "" if cpu_only else visible(k)
        tasks.append(dict(name=f"ps{k}" if num_ps > 1 else "ps", job="ps", index=k,
                          env={"CUDA_VISIBLE_DEVICES": vis},
                          argv=[python, MAIN, *common, "--job_name", "ps", f"--task_index={k}"]))
    for i in range(num_workers):
        vis = "" if cpu_only else visible(i)                          # ID=$((i % num_gpus)), reference :20
        tasks.append(dict(name=f"worker{i}", job="worker", index=i, env={"CUDA_VISIBLE_DEVICES": vis},
                          argv=[python, MAIN, *common, "--job_name", "worker", "--task_index", str(i)]))
    return tasks


# ---- tmux mode ------------------------------------------------------------------------------------
def launch_tmux(session: str, tasks: List[Dict], attach: bool = True, dry_run: bool = False) -> List[List[str]]:
    cmds: List[List[str]] = [["tmux", "kill-session", "-t", session]]
    first = True
    for t in tasks:
        if first:
            cmds.append(["tmux", "new-session", "-s", session, "-n", t["name"], "-d", "bash"])
            first = False
        else:
            cmds.append(["tmux", "new-window", "-t", session, "-n", t["name"], "-d", "bash"])
    for t in tasks:
        env = " ".join(f"{k}={v}" for k, v in t["env"].items())
        line = f"cd {shlex.quote(REPO)} && {env} " + " ".join(shlex.quote(a) for a in t["argv"])
        cmds.append(["tmux", "send-keys", "-t", f"{session}:{t['name']}", line, "Enter"])
    if attach:
        cmds.append(["tmux", "a", "-t", session])
    if dry_run:
        return cmds
    if shutil.which("tmux") is None:
        raise RuntimeError("tmux not found; use --mode headless")
    for i, c in enumerate(cmds):
        if c[1] == "send-keys" and cmds[i - 1][1] != "send-keys":
            time.sleep(1)                       # let the shells come up (reference: sleep 1)
        subprocess.run(c, check=False, stderr=subprocess.DEVNULL if c[1] == "kill-session" else None)
    return cmds


# ---- headless supervising mode ----------------------------------------------------------------------
class HeadlessCluster:
    def __init__(self, tasks: List[Dict], log_dir: Optional[str] = None, echo: bool = True):
        self.tasks, self.echo, self.log_dir = tasks, echo, log_dir
        self.procs: Dict[str, subprocess.Popen] = {}
        self._pumps: List[threading.Thread] = []
        self.lines: Dict[str, List[str]] = {t["name"]: [] for t in tasks}

    def start(self) -> None:
        if self.log_dir:
            os.makedirs(self.log_dir, exist_ok=True)
        for t in self.tasks:
            env = dict(os.environ)
            env.update(t["env"])
            env["PYTHONUNBUFFERED"] = "1"
            p = subprocess.Popen(t["argv"], cwd=REPO, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                                 text=True, start_new_session=True)
            self.procs[t["name"]] = p
            th = threading.Thread(target=self._pump, args=(t["name"], p), daemon=True)
            th.start()
            self._pumps.append(th)

    def _pump(self, name: str, p: subprocess.Popen) -> None:
        logf = open(os.path.join(self.log_dir, name + ".log"), "w") if self.log_dir else None
        for line in p.stdout:
            self.lines[name].append(line.rstrip("\n"))
            if self.echo:
                sys.stdout.write(f"[{name}] {line}")
                sys.stdout.flush()
            if logf:
                logf.write(line)
                logf.flush()
        if logf:
            logf.close()

    def signal_all(self, sig: int, jobs: Sequence[str] = ("ps", "worker")) -> None:
        for t in self.tasks:
            p = self.procs.get(t["name"])
            if p is not None and p.poll() is None and t["job"] in jobs:
                try:
                    os.killpg(p.pid, sig)       # the exact process group this launcher created
                except ProcessLookupError:
                    pass

    def wait(self, timeout: Optional[float] = None, ps_grace_s: float = 10.0) -> Dict[str, int]:
        """Wait for every worker, then stop the ps tasks; returns exit codes by task name."""
        deadline = None if timeout is None else time.time() + timeout
        workers = [t["name"] for t in self.tasks if t["job"] == "worker"]
        for name in workers:
            p = self.procs[name]
            remaining = None if deadline is None else max(0.1, deadline - time.time())
            try:
                p.wait(timeout=remaining)
            except subprocess.TimeoutExpired:
                self.signal_all(signal.SIGTERM)
                break
        # gloo ps tasks exit on their own once all workers said goodbye; others need the signal
        t_end = time.time() + ps_grace_s
        for t in self.tasks:
            if t["job"] != "ps":
                continue
            p = self.procs[t["name"]]
            try:
                p.wait(timeout=max(0.1, min(2.0, t_end - time.time())))
            except subprocess.TimeoutExpired:
                self.signal_all(signal.SIGTERM, jobs=("ps",))
        codes = {}
        for name, p in self.procs.items():
            try:
                codes[name] = p.wait(timeout=ps_grace_s)
            except subprocess.TimeoutExpired:
                os.killpg(p.pid, signal.SIGKILL)
                codes[name] = p.wait()
        for th in self._pumps:
            th.join(timeout=2)
        return codes


def launch_headless(tasks: List[Dict], timeout: Optional[float] = None, log_dir: Optional[str] = None) -> int:
    cluster = HeadlessCluster(tasks, log_dir=log_dir)
    cluster.start()

    def forward(signum, frame):
        cluster.signal_all(signum)
    old = {s: signal.signal(s, forward) for s in (signal.SIGINT, signal.SIGTERM, signal.SIGHUP)}
    try:
        codes = cluster.wait(timeout)
    finally:
        for s, h in old.items():
            signal.signal(s, h)
    worker_codes = [c for n, c in codes.items() if n.startswith("worker")]
    # ps tasks are stopped with SIGTERM by design (-15 / 143), which is not a failure
    worst = max((abs(c) for c in worker_codes), default=0)
    print(f"[launcher] exit codes: {codes}", flush=True)
    return worst


def main(argv: Optional[Sequence[str]] = None) -> int:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--name", default="sess", help="tmux session name")
    ap.add_argument("--num_workers", type=int, default=2)
    ap.add_argument("--num_gpus", type=int, default=1)
    ap.add_argument("--gpu_id", default="0", help="the GPU_ID array, comma or space separated")
    ap.add_argument("--num_ps", type=int, default=1)
    ap.add_argument("--device", default="auto", choices=["auto", "cuda", "cpu"])
    ap.add_argument("--mode", default="tmux", choices=["tmux", "headless", "print"])
    ap.add_argument("--timeout", type=float, default=None)
    ap.add_argument("--log_dir", default=None)
    ap.add_argument("--no-attach", action="store_true")
    args, extra = ap.parse_known_args(argv)
    extra = [a for a in extra if a != "--"]      # `launcher ... -- --training_steps 200`: the separator is not a flag
    gpu_ids = [g for g in args.gpu_id.replace(",", " ").split() if g]
    tasks = task_commands(args.num_workers, args.num_gpus, gpu_ids, args.num_ps, args.device, extra)
    if args.mode == "print":
        for c in launch_tmux(args.name, tasks, attach=not args.no_attach, dry_run=True):
            print(" ".join(shlex.quote(a) for a in c))
        return 0
    if args.mode == "tmux":
        launch_tmux(args.name, tasks, attach=not args.no_attach)
        return 0
    return launch_headless(tasks, timeout=args.timeout, log_dir=args.log_dir)


if __name__ == "__main__":
    sys.exit(main())
