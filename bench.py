#!/usr/bin/env python
"""Headline benchmark: MNIST MLP (784-100-10 sigmoid/softmax, batch 100 per worker, SGD lr 0.001)
asynchronous parameter-server training, examples/sec, device-timed, max over ranks.

    python bench.py --gpus N --steps K --warmup W                 # this framework
    python bench.py --impl reference ...                          # the TF1 reference (unavailable here)
    python bench.py --impl nccl-ps ...                            # torch.distributed NCCL ps baseline
    python bench.py --config wide ...                             # 784-8192-8192-10, batch 4096, bf16
    python bench.py --config bw --gpus N ...                      # ps push/pull bandwidth sweep 4 KB - 1 GB vs NCCL

For N>1 the driver launches this file under torchrun (one rank per GPU).  One JSON line is
printed by rank 0.  See DESIGN.md "Benchmark methodology" for what is inside the timed region.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "MNIST MLP examples/sec (device-timed, max over ranks)"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "nccl-ps"])
    ap.add_argument("--config", default="mnist", choices=["mnist", "wide", "bw"])
    ap.add_argument("--bw-max-mb", dest="bw_max_mb", type=int, default=1024, help="--config bw: largest shard in MB")
    ap.add_argument("--mode", default="async", choices=["async", "sync"])
    ap.add_argument("--batch", type=int, default=0, help="per-worker batch (default: 100 mnist / 4096 wide)")
    ap.add_argument("--steps-per-launch", type=int, default=0,
                    help="training steps fused into one persistent-kernel launch (0 = auto)")
    ap.add_argument("--engine", default="auto", choices=["auto", "fused", "fused_tc", "layerwise", "tc"])
    ap.add_argument("--e2e-steps", type=int, default=0, help="steps of the host-fed end-to-end measurement (0 = --steps)")
    ap.add_argument("--e2e-steps-per-launch", type=int, default=8,
                    help="host-fed steps that share one kernel launch in the end-to-end measurement (1..16)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--sync-mode", dest="sync_mode", default="allreduce_ll", choices=["allreduce_ll", "allreduce", "reduce_scatter"],
                    help="--mode sync: form of the in-kernel exchange")
    ap.add_argument("--repeats", type=int, default=25, help="timed samples of K steps each (median is reported)")
    ap.add_argument("--no-baseline", dest="no_baseline", action="store_true",
                    help="skip the in-run NCCL baseline (vs_baseline stays null)")
    ap.add_argument("--no-entry", dest="no_entry", action="store_true",
                    help="skip the launcher-path (main.py) measurement")
    ap.add_argument("--entry-steps", dest="entry_steps", type=int, default=0,
                    help="training steps per worker of the launcher-path measurement (0 = 300000)")
    ap.add_argument("--entry-multi", dest="entry_multi", action="store_true",
                    help="with 4 GPUs also run the reference's 16 workers / 4 GPUs topology")
    return ap.parse_args()


# ----------------------------------------------------------------------------------------------
# clocks: sampled with NVML while the timed region runs
# ----------------------------------------------------------------------------------------------
class ClockSampler:
    REASONS = {
        0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
        0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
        0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting",
    }

    def __init__(self, index: int, period_s: float = 0.005):
        self.samples, self.reason_bits, self.period = [], 0, period_s
        self._stop = threading.Event()
        self._thread = None
        self.max_mhz = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nv = pynvml
            # NVML indexes physical devices; honour CUDA_VISIBLE_DEVICES when it is a plain index list
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            phys = index
            if vis and all(p.strip().isdigit() for p in vis.split(",")):
                ids = [int(p) for p in vis.split(",")]
                if index < len(ids):
                    phys = ids[index]
            self._h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self._nv = None

    def _sample(self):
        nv = self._nv
        try:
            self.samples.append(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM))
            self.reason_bits |= int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._h))
        except Exception:
            pass

    def _run(self):
        while not self._stop.is_set():
            self._sample()
            self._stop.wait(self.period)

    def start(self):
        if self._nv is None:
            return
        self._sample()
        self._thread = threading.Thread(target=self._run, daemon=True)
        self._thread.start()

    def stop(self):
        if self._nv is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml_unavailable"]}
        self._stop.set()
        if self._thread:
            self._thread.join()
        self._sample()
        s = sorted(self.samples)
        med = s[len(s) // 2] if s else None
        reasons = [name for bit, name in self.REASONS.items() if self.reason_bits & bit and name != "gpu_idle"]
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": reasons, "samples": len(s)}


def _reference_unavailable():
    why = ("TensorFlow 1.x is not installable here: /root/reference ships no setup.py/wheels (pip install of the "
           "source tree fails), no network, Python 3.12 (TF1 needs <=3.7), and TF 1.x predates CUDA 12.8/sm_100")
    if int(os.environ.get("RANK", "0")) == 0:       # one line per job, also when launched under torchrun
        print(json.dumps({"impl": "reference", "unavailable": why}), flush=True)


def main():
    args = parse_args()
    if args.impl == "reference":
        _reference_unavailable()
        return 0

    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        print(json.dumps({"error": "no CUDA device; bench.py needs a B200", "metric": METRIC}))
        return 1
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    if args.impl == "nccl-ps":
        from baseline import nccl_ps
        return nccl_ps.run(args, rank, world, dev)

    from distributedtensorflowexample_b200 import benchlib
    if args.config == "wide":
        return benchlib.run_wide(args, rank, world, dev, ClockSampler)
    if args.config == "bw":
        from distributedtensorflowexample_b200 import bench_bw
        return bench_bw.run(args, rank, world, dev)
    return benchlib.run_mnist(args, rank, world, dev, ClockSampler)


if __name__ == "__main__":
    sys.exit(main() or 0)
