"""Time-to-target test accuracy THROUGH THE ENTRY POINT (main.py started by the headless launcher), at N workers
one per GPU: asynchronous (pull one step ahead, and DTFE_MLP2_NO_AHEAD=1 = the reference's pull-then-step order)
and --sync.  Real MNIST files that ship with the reference (7500/500/2000 split of t10k: train-images is absent),
batch 100 per worker, SGD lr 0.1, target 88 % on the 2000 held-out images, evaluated by whichever worker observes a
global step that is a multiple of --eval-every (reference worker.py:150-154).

Clock: wall_time of the `local/test_accuracy` event that first reaches the target minus wall_time of the earliest
`local/loss` event of any worker (both written by the workers themselves); process start-up is not in it.

    python scripts/tta_launcher.py --gpus 1 2 4 8
"""
import argparse, json, os, socket, sys, tempfile, shutil
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributedtensorflowexample_b200.launch import launcher
from distributedtensorflowexample_b200.utils.events_reader import read_events

ap = argparse.ArgumentParser()
ap.add_argument("--gpus", type=int, nargs="+", default=[1])
ap.add_argument("--target", type=float, default=0.88)
ap.add_argument("--lr", type=float, default=0.1)
ap.add_argument("--eval-every", type=int, default=500)
ap.add_argument("--max-steps", type=int, default=60000)
ap.add_argument("--repeats", type=int, default=3)
ap.add_argument("--modes", nargs="+", default=["async", "async_no_ahead", "sync"])
args = ap.parse_args()


def run(n, mode, seed):
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    logdir = tempfile.mkdtemp(prefix="dtfe_tta_")
    extra = ["--logdir", logdir + "/log", "--data", "mnist", "--seed", str(seed), "--training_steps", str(args.max_steps),
             "--learning_rate", str(args.lr), "--eval_every", str(args.eval_every), "--recovery_wait_secs", "1",
             "--base_port", str(port if port + 40 < 65000 else 24000), "--save_model_secs", "600"]
    num_ps = 1
    if mode == "sync":
        extra.append("--sync"); num_ps = n
    env_key = "DTFE_MLP2_NO_AHEAD"
    if mode == "async_no_ahead":
        os.environ[env_key] = "1"
    else:
        os.environ.pop(env_key, None)
    ids = [str(i) for i in range(n)]
    tasks = launcher.task_commands(n, n, ids, num_ps=num_ps, device="cuda", extra=extra)
    cl = launcher.HeadlessCluster(tasks, echo=False)
    cl.start()
    codes = cl.wait(timeout=300)
    os.environ.pop(env_key, None)
    first_loss, hits, best = None, [], 0.0
    for w in range(n):
        for e in read_events(f"{logdir}/log_{w}"):
            for tag, v in e.scalars:
                if tag == "local/loss" and (first_loss is None or e.wall_time < first_loss):
                    first_loss = e.wall_time
                if tag == "local/test_accuracy":
                    best = max(best, v)
                    if v >= args.target:
                        hits.append((e.wall_time, e.step, v))
    shutil.rmtree(logdir, ignore_errors=True)
    out = {"workers": n, "mode": mode, "seed": seed, "target": args.target, "lr": args.lr, "eval_every": args.eval_every,
           "ok": all(codes.get(f"worker{i}") == 0 for i in range(n)), "best_accuracy": round(best, 4)}
    if hits and first_loss is not None:
        t, step, acc = min(hits)
        out.update(reached=True, seconds=round(t - first_loss, 4), global_step=step, test_accuracy=round(acc, 4))
    else:
        out.update(reached=False)
    return out


for n in args.gpus:
    for mode in args.modes:
        for rep in range(args.repeats):
            print(json.dumps(run(n, mode, seed=rep)), flush=True)
