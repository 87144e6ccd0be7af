"""Builds the native library in-tree: ``csrc/*.cu`` -> ``dla_b200/ops/libdla_b200.so``.

nvcc cross-compiles sm_100a without a GPU, so this runs on the CPU build box; the ``.so`` is
git-ignored but travels with the working tree to the GPU box.  ``python -m dla_b200.ops.build``.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path
from typing import List

ROOT = Path(__file__).resolve().parents[2]
CSRC = ROOT / "csrc"
OUT_DIR = Path(__file__).resolve().parent
LIB = OUT_DIR / "libdla_b200.so"
OBJ_DIR = ROOT / "build" / "obj"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr", "-Xptxas", "-v",
]


# extra -D switches for diagnostic builds, e.g. DLA_BUILD_DEFINES="-DDLA_PEER_DEBUG" (signal-pad sequence assertions)
NVCC_FLAGS += [f for f in os.environ.get("DLA_BUILD_DEFINES", "").split() if f.startswith("-D")]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found")


def _sources() -> List[Path]:
    return sorted(CSRC.glob("*.cu"))


def _digest() -> str:
    h = hashlib.sha256()
    for p in sorted(list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cuh")) + list((ROOT / "include").glob("*.h"))):
        h.update(p.name.encode())
        h.update(p.read_bytes())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def is_current() -> bool:
    stamp = OUT_DIR / ".libdla_b200.stamp"
    return LIB.exists() and stamp.exists() and stamp.read_text().strip() == _digest()


def build(force: bool = False, verbose: bool = False) -> Path:
    if not force and is_current():
        return LIB
    nvcc = _nvcc()
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    log_dir = ROOT / "build" / "logs"
    log_dir.mkdir(parents=True, exist_ok=True)

    def compile_one(src: Path) -> Path:
        obj = OBJ_DIR / (src.stem + ".o")
        cmd = [nvcc, *NVCC_FLAGS, "-I", str(ROOT / "include"), "-c", str(src), "-o", str(obj)]
        res = subprocess.run(cmd, capture_output=True, text=True)
        (log_dir / (src.stem + ".ptxas.log")).write_text(res.stdout + res.stderr)
        if res.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src.name}:\n{res.stdout}\n{res.stderr}")
        if verbose:
            print(res.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 2)) as ex:
        objs = list(ex.map(compile_one, _sources()))
    tmp = LIB.with_suffix(".so.tmp")
    cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", str(tmp), *map(str, objs)]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"link failed:\n{res.stdout}\n{res.stderr}")
    os.replace(tmp, LIB)
    (OUT_DIR / ".libdla_b200.stamp").write_text(_digest())
    return LIB


def main() -> None:
    """``python -m dla_b200.ops.build [--force] [-v]`` / the ``dla-build`` console script."""
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))


if __name__ == "__main__":
    main()
