"""In-tree native build: two shared libraries, no torch headers, seconds to compile.

* ``_lib/libdtfe_cuda.so``  - every CUDA kernel, compiled for sm_100a only
  (``-gencode arch=compute_100a,code=sm_100a -lineinfo``), C ABI (``include/dtfe_b200.h``).
* ``_lib/libdtfe_host.so``  - the C++ runtime pieces that need no GPU: crc32c, SSTable,
  TF1 tensor-bundle checkpoints, tfevents, idx/gzip reader (the rendezvous is torch's TCPStore, train/server.py).

The reference ships no native code at all; these replace the parts of the TF1
runtime it leans on.  Both libraries are loaded with ctypes, so the built ``.so`` files
travel with the source tree and are visible in ``/proc/self/maps`` of every process that
uses them.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent
REPO_DIR = PKG_DIR.parent
CSRC = REPO_DIR / "csrc"
LIB_DIR = PKG_DIR / "_lib"
CUDA_LIB = LIB_DIR / "libdtfe_cuda.so"
HOST_LIB = LIB_DIR / "libdtfe_host.so"

NVCC_ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found (looked at $NVCC, PATH, /usr/local/cuda/bin)")


def _newest_mtime(paths) -> float:
    return max((p.stat().st_mtime for p in paths), default=0.0)


def _stale(target: Path, sources) -> bool:
    return (not target.exists()) or target.stat().st_mtime < _newest_mtime(sources)


def _run(cmd, verbose: bool) -> None:
    if verbose:
        print("[dtfe build]", " ".join(str(c) for c in cmd), flush=True)
    proc = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if proc.returncode != 0:
        sys.stderr.write(proc.stdout)
        raise RuntimeError(f"native build failed: {' '.join(str(c) for c in cmd)}")
    if verbose and proc.stdout.strip():
        print(proc.stdout)


def cuda_sources():
    return sorted((CSRC / "cuda").glob("*.cu"))


def build_cuda(force: bool = False, verbose: bool = False, ptxas_verbose: bool = False) -> Path:
    srcs = cuda_sources()
    deps = srcs + sorted((CSRC / "cuda").glob("*.cuh")) + sorted((REPO_DIR / "include").glob("*.h"))
    LIB_DIR.mkdir(exist_ok=True)
    obj_dir = REPO_DIR / "build" / "cuda"
    obj_dir.mkdir(parents=True, exist_ok=True)
    hdr_mtime = _newest_mtime([d for d in deps if d.suffix != ".cu"])
    objs = []
    procs = []
    for src in srcs:
        obj = obj_dir / (src.stem + ".o")
        objs.append(obj)
        if force or not obj.exists() or obj.stat().st_mtime < max(src.stat().st_mtime, hdr_mtime):
            cmd = [_nvcc(), "-O3", "-std=c++17", *NVCC_ARCH, "-lineinfo", "-Xcompiler", "-fPIC,-fvisibility=hidden",
                   "-I", str(REPO_DIR / "include"), "-c", str(src), "-o", str(obj)]
            if ptxas_verbose:
                cmd += ["-Xptxas", "-v"]
            if verbose:
                print("[dtfe build]", " ".join(cmd), flush=True)
            procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = False
    for cmd, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            sys.stderr.write(out)
            failed = True
        elif (verbose or ptxas_verbose) and out.strip():
            print(out)
    if failed:
        raise RuntimeError("nvcc failed (see output above)")
    if force or procs or _stale(CUDA_LIB, objs):
        _run([_nvcc(), "-shared", *NVCC_ARCH, "-o", str(CUDA_LIB), *map(str, objs), "-cudart", "static"], verbose)
    return CUDA_LIB


def host_sources():
    return sorted((CSRC / "host").glob("*.cc"))


def build_host(force: bool = False, verbose: bool = False) -> Path:
    srcs = host_sources()
    deps = srcs + sorted((CSRC / "host").glob("*.h"))
    LIB_DIR.mkdir(exist_ok=True)
    if force or _stale(HOST_LIB, deps):
        _run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-fvisibility=hidden", "-Wall",
              "-o", str(HOST_LIB), *map(str, srcs), "-lz", "-lpthread"], verbose)
    return HOST_LIB


def build_all(force: bool = False, verbose: bool = False) -> None:
    build_host(force, verbose)
    build_cuda(force, verbose)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv, verbose=True)
