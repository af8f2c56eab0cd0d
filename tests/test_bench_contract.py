"""CPU checks of the bench.py driver contract that do not need a GPU."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(args, env=None):
    e = dict(os.environ)
    e.update(env or {})
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True,
                          env=e, timeout=300, cwd=ROOT)


def test_reference_arm_reports_unavailable_once_per_job():
    r = _run(["--impl", "reference", "--gpus", "1", "--steps", "5", "--warmup", "3"])
    assert r.returncode == 0
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and "unavailable" in d and "\n" not in d["unavailable"]
    # under torchrun every rank runs the script; only rank 0 may print
    r1 = _run(["--impl", "reference", "--gpus", "8"], env={"RANK": "3", "WORLD_SIZE": "8", "LOCAL_RANK": "3"})
    assert r1.returncode == 0 and r1.stdout.strip() == ""


def test_own_arm_without_gpu_fails_loudly():
    import torch
    if torch.cuda.is_available():
        return
    r = _run(["--steps", "3", "--warmup", "3"])
    assert r.returncode != 0 and "no CUDA device" in r.stdout


def test_graft_entry_build_produces_loadable_libraries():
    sys.path.insert(0, ROOT)
    import __graft_entry__ as g
    g.build()
    from distributedtensorflowexample_b200 import _build
    assert _build.CUDA_LIB.exists() and _build.HOST_LIB.exists()
    # every entry point declared in the public header is exported by the CUDA library
    import ctypes, re
    lib = ctypes.CDLL(str(_build.CUDA_LIB))
    hdr = open(os.path.join(ROOT, "include", "dtfe_b200.h")).read()
    names = set(re.findall(r"\b(dtfe_[a-z0-9_]+)\s*\(", hdr))
    assert len(names) > 25
    for n in sorted(names):
        assert hasattr(lib, n), f"{n} is declared in include/dtfe_b200.h but not exported"


def test_ctypes_struct_mirrors_match_the_c_layout():
    """The ctypes mirrors of the launch structs must have the C compiler's size (field order/padding drift
    between include/dtfe_b200.h and the bindings would otherwise corrupt arguments silently)."""
    import ctypes
    from distributedtensorflowexample_b200.ops import cuda_ops, gemm_tc
    sizes = (ctypes.c_size_t * 3)()
    assert cuda_ops.lib().dtfe_struct_sizes(sizes, 3) == 3
    assert ctypes.sizeof(cuda_ops._Mlp2Launch) == sizes[0]
    assert ctypes.sizeof(cuda_ops.SyncExchange) == sizes[1]
    assert ctypes.sizeof(gemm_tc._GemmTc) == sizes[2]
