"""In-tree build of the native libraries.

Two shared objects are produced next to this file:

* ``libb200zero.so``      — every ``csrc/*.cu`` compiled with nvcc for sm_100a only
  (``-gencode arch=compute_100a,code=sm_100a -lineinfo``).
* ``libb200zero_host.so`` — ``csrc/host/*.cpp`` (host Adam, checkpoint writer), plain g++.

The build is incremental (object files under ``build/``, keyed by source mtime and
flags) and parallel across translation units. nvcc cross-compiles without a GPU,
so this runs on the CPU builder box; the resulting ``.so`` files travel to the GPU
box with the repo snapshot.

The reference has no build step at all (no native code: SURVEY.md §2.2).
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

PKG_OPS = Path(__file__).resolve().parent
REPO = PKG_OPS.parent.parent
CSRC = REPO / "csrc"
BUILD = REPO / "build"
LIB_CUDA = PKG_OPS / "libb200zero.so"
LIB_HOST = PKG_OPS / "libb200zero_host.so"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "-Xcompiler", "-fvisibility=default",
]
CXX_FLAGS = ["-O3", "-std=c++17", "-fPIC", "-march=x86-64-v3", "-pthread", "-fopenmp-simd"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found; cannot build libb200zero.so")


def _stamp(src: Path, flags: List[str]) -> str:
    h = hashlib.sha1()
    h.update(src.read_bytes())
    for hdr in sorted(list(CSRC.glob("*.cuh")) + list((REPO / "include").glob("*.h"))):
        h.update(hdr.read_bytes())
    h.update(" ".join(flags).encode())
    return h.hexdigest()


def _compile_one(compiler: str, flags: List[str], src: Path, obj: Path, verbose: bool) -> None:
    stamp_file = obj.with_suffix(obj.suffix + ".stamp")
    stamp = _stamp(src, flags)
    if obj.exists() and stamp_file.exists() and stamp_file.read_text() == stamp:
        return
    cmd = [compiler, *flags, "-c", str(src), "-o", str(obj)]
    if verbose:
        print("[build]", " ".join(cmd), flush=True)
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"compile failed: {' '.join(cmd)}\n{res.stdout}\n{res.stderr}")
    stamp_file.write_text(stamp)


def build_cuda(verbose: bool = False, force: bool = False) -> Path:
    srcs = sorted(CSRC.glob("*.cu"))
    if not srcs:
        raise RuntimeError(f"no CUDA sources under {CSRC}")
    BUILD.mkdir(exist_ok=True)
    nvcc = _nvcc()
    objs = [BUILD / (s.stem + ".o") for s in srcs]
    if force:
        for o in objs:
            o.unlink(missing_ok=True)
    newest_obj_before = max((o.stat().st_mtime for o in objs if o.exists()), default=0.0)
    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 2)) as ex:
        list(ex.map(lambda so: _compile_one(nvcc, NVCC_FLAGS, so[0], so[1], verbose), zip(srcs, objs)))
    newest_obj = max(o.stat().st_mtime for o in objs)
    if (not LIB_CUDA.exists()) or newest_obj > LIB_CUDA.stat().st_mtime or newest_obj != newest_obj_before:
        cmd = [nvcc, "-shared", "-o", str(LIB_CUDA), *map(str, objs)]  # cudart linked statically (nvcc default): immune to torch/toolkit skew
        if verbose:
            print("[build]", " ".join(cmd), flush=True)
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"link failed: {res.stdout}\n{res.stderr}")
    return LIB_CUDA


def build_host(verbose: bool = False, force: bool = False) -> Path:
    srcs = sorted((CSRC / "host").glob("*.cpp"))
    if not srcs:
        return LIB_HOST
    BUILD.mkdir(exist_ok=True)
    cxx = os.environ.get("CXX") or shutil.which("g++") or "g++"
    objs = [BUILD / ("host_" + s.stem + ".o") for s in srcs]
    if force:
        for o in objs:
            o.unlink(missing_ok=True)
    with ThreadPoolExecutor(max_workers=4) as ex:
        list(ex.map(lambda so: _compile_one(cxx, CXX_FLAGS, so[0], so[1], verbose), zip(srcs, objs)))
    newest_obj = max(o.stat().st_mtime for o in objs)
    if (not LIB_HOST.exists()) or newest_obj > LIB_HOST.stat().st_mtime:
        cmd = [cxx, "-shared", "-o", str(LIB_HOST), *map(str, objs), "-pthread"]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"host link failed: {res.stdout}\n{res.stderr}")
    return LIB_HOST


def build_all(verbose: bool = False, force: bool = False) -> None:
    build_cuda(verbose=verbose, force=force)
    build_host(verbose=verbose, force=force)


if __name__ == "__main__":
    build_all(verbose=True, force="--force" in sys.argv)
    print("built:", LIB_CUDA, LIB_HOST if LIB_HOST.exists() else "(no host lib)")
