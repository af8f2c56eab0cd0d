"""Per-process GPU memory budgeting (reference main.py:57-64,73).

``fraction = 0.9 / ceil(num_workers / num_gpus)``: 2 workers on 1 GPU -> 0.45, 16 on 4 -> 0.225.
TF1 applies it as ``per_process_gpu_memory_fraction``; here it caps torch's caching allocator."""
from __future__ import annotations

import math


def per_process_gpu_memory_fraction(num_workers: int, num_gpus: int) -> float:
    if num_gpus < 1 or num_workers < 1:
        raise ValueError("num_workers and num_gpus must be >= 1")
    return 0.9 / math.ceil(float(num_workers) / float(num_gpus))


def banner(fraction: float) -> str:
    rule = "-" * 100
    return f"{rule}\nPer-process GPU memory fraction: {fraction}\n{rule}"


def apply_fraction(fraction: float, device=None) -> bool:
    """Returns True if a CUDA device was configured."""
    import torch
    if not torch.cuda.is_available():
        return False
    torch.cuda.set_per_process_memory_fraction(min(max(fraction, 0.0), 1.0), device)
    return True
