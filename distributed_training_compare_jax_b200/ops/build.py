"""Build the sm_100a kernel library in-tree with nvcc (no torch headers needed, so it takes seconds).

    python -m distributed_training_compare_jax_b200.ops.build [--force]

Every ``csrc/*.cu`` is compiled with ``-gencode arch=compute_100a,code=sm_100a -lineinfo`` into
``build/obj/*.o`` and linked into ``distributed_training_compare_jax_b200/lib/libdtc_b200.so``.  The ``.so`` is
git-ignored but travels to the GPU box with the repo snapshot.  nvcc cross-compiles without a GPU.
"""
from __future__ import annotations

import concurrent.futures as cf
import hashlib
import os
import shutil
import subprocess
import sys
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent.parent
REPO_DIR = PKG_DIR.parent
CSRC = PKG_DIR / "csrc"
LIB_DIR = PKG_DIR / "lib"
LIB_PATH = LIB_DIR / "libdtc_b200.so"
OBJ_DIR = REPO_DIR / "build" / "obj"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
    "-I", str(REPO_DIR / "include"),
    "-I", str(CSRC),
]
# debug facilities compiled in on request, e.g. DTC_NVCC_DEFINES=DTC_GEMM_TRACE (part of the digest: set it wherever the
# library is loaded as well, or the loader rebuilds without it)
NVCC_FLAGS += [f"-D{d}" for d in os.environ.get("DTC_NVCC_DEFINES", "").split(",") if d]


def find_nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found (set NVCC=/path/to/nvcc)")


def _sources() -> list[Path]:
    return sorted(CSRC.glob("*.cu"))


def _digest() -> str:
    h = hashlib.sha256()
    for p in sorted(list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.h"))
                    + [REPO_DIR / "include" / "dtc_b200.h"]):
        h.update(p.name.encode())
        h.update(p.read_bytes())
    # flags without the absolute include paths: the same tree must digest identically wherever it is checked out
    # (the GPU box runs a copy of the repo under a scratch path)
    h.update(" ".join(f for f in NVCC_FLAGS if not os.path.isabs(f)).encode())
    return h.hexdigest()


def is_fresh() -> bool:
    stamp = LIB_DIR / "libdtc_b200.digest"
    return LIB_PATH.exists() and stamp.exists() and stamp.read_text().strip() == _digest()


def build(force: bool = False, verbose: bool = True) -> Path:
    if not force and is_fresh():
        return LIB_PATH
    nvcc = find_nvcc()
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    # several ranks may get here at once: one builds, the others wait for the lock and find a fresh library
    import fcntl
    lock = open(LIB_DIR / ".build.lock", "w")
    fcntl.flock(lock, fcntl.LOCK_EX)
    try:
        if not force and is_fresh():
            return LIB_PATH
        return _build_locked(nvcc, verbose)
    finally:
        fcntl.flock(lock, fcntl.LOCK_UN)
        lock.close()


def _build_locked(nvcc: str, verbose: bool) -> Path:

    def compile_one(src: Path) -> Path:
        obj = OBJ_DIR / (src.stem + ".o")
        cmd = [nvcc, *NVCC_FLAGS, "-c", str(src), "-o", str(obj)]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src.name}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            print(f"[dtc build] {src.name} ok", flush=True)
        return obj

    with cf.ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 4)) as ex:
        objs = list(ex.map(compile_one, _sources()))
    cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", str(LIB_PATH), *map(str, objs)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    (LIB_DIR / "libdtc_b200.digest").write_text(_digest())
    if verbose:
        print(f"[dtc build] linked {LIB_PATH}", flush=True)
    return LIB_PATH


if __name__ == "__main__":
    build(force="--force" in sys.argv)
