"""In-tree build of libb2flat.so (nvcc, sm_100a only).

`python -m nv_genai_b200.build [--force] [-v]` compiles every .cu under csrc/ to an object
(in parallel) and links them into nv_genai_b200/lib/libb2flat.so.  The shared library has
no torch / Python dependency: it is the C-ABI drop-in declared in include/b2_flat.h.
"""
from __future__ import annotations

import concurrent.futures
import glob
import os
import shutil
import subprocess
import sys

_PKG = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_PKG, "csrc")
LIBDIR = os.path.join(_PKG, "lib")
OBJDIR = os.path.join(LIBDIR, "obj")
SO_PATH = os.path.join(LIBDIR, "libb2flat.so")

ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_FLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
              "--expt-relaxed-constexpr", "-Xptxas", "-v"]


def _nvcc() -> str:
    cand = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(cand):
        raise RuntimeError("nvcc not found; set NVCC or add /usr/local/cuda/bin to PATH")
    return cand


def _sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def _deps_mtime() -> float:
    files = _sources() + glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(CSRC, "*.h"))
    files.append(os.path.join(os.path.dirname(_PKG), "include", "b2_flat.h"))
    return max(os.path.getmtime(f) for f in files)


def needs_build() -> bool:
    return (not os.path.exists(SO_PATH)) or os.path.getmtime(SO_PATH) < _deps_mtime()


def _compile_one(nvcc: str, src: str, verbose: bool) -> str:
    obj = os.path.join(OBJDIR, os.path.basename(src)[:-3] + ".o")
    cmd = [nvcc] + ARCH_FLAGS + NVCC_FLAGS + ["-c", src, "-o", obj]
    res = subprocess.run(cmd, capture_output=True, text=True)
    log = os.path.join(OBJDIR, os.path.basename(src)[:-3] + ".ptxas.log")
    with open(log, "w") as f:
        f.write(res.stdout + res.stderr)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed for %s:\n%s" % (src, (res.stdout + res.stderr)[-6000:]))
    if verbose:
        print("compiled", os.path.basename(src))
    return obj


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return SO_PATH
    os.makedirs(OBJDIR, exist_ok=True)
    nvcc = _nvcc()
    srcs = _sources()
    with concurrent.futures.ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(lambda s: _compile_one(nvcc, s, verbose), srcs))
    tmp = SO_PATH + ".tmp"
    cmd = [nvcc] + ARCH_FLAGS + ["-shared", "-o", tmp] + objs + ["-Xcompiler", "-fvisibility=hidden"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("link failed:\n" + res.stdout + res.stderr)
    os.replace(tmp, SO_PATH)
    if verbose:
        print("linked", SO_PATH)
    return SO_PATH


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
