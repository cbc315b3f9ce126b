"""In-tree build of libserenade_b200.so (sm_100a only, nvcc; no JIT cache, no torch extension).

    python -m serenade_b200.build [--force]
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG, "csrc")
LIB = os.path.join(PKG, "libserenade_b200.so")
STAMP = os.path.join(PKG, "build", "stamp.txt")
SOURCES = ["api.cu", "engine.cu", "noalloc.cu"]
NVCC_FLAGS = ["-std=c++17", "-O3", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found (needed to build libserenade_b200.so for sm_100a)")


def _source_hash() -> str:
    h = hashlib.sha256()
    files = sorted(os.listdir(CSRC)) + ["../../include/serenade_b200.h"]
    for f in files:
        p = os.path.join(CSRC, f)
        if os.path.isfile(p):
            h.update(f.encode())
            h.update(open(p, "rb").read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def build_library(force: bool = False, verbose: bool = False, variant: str = "", defines=()) -> str:
    """Builds libserenade_b200.so.  `variant` / `defines` build an experiment copy libserenade_b200_<variant>.so
    with extra -D flags (selected at run time with SRN_LIB_VARIANT=<variant>; A/B timing on one GPU box)."""
    if variant:
        return _build(os.path.join(PKG, f"libserenade_b200_{variant}.so"), os.path.join(PKG, "build", variant),
                      [f"-D{d}" for d in defines], True, verbose)
    want = _source_hash()
    if not force and os.path.exists(LIB) and os.path.exists(STAMP) and open(STAMP).read().strip() == want:
        return LIB
    _build(LIB, os.path.join(PKG, "build"), [], force, verbose)
    with open(STAMP, "w") as f:
        f.write(want)
    return LIB


def _build(lib: str, objdir: str, extra, force: bool, verbose: bool) -> str:
    nvcc = _nvcc()
    os.makedirs(objdir, exist_ok=True)
    objs = []
    procs = []
    for src in SOURCES:
        obj = os.path.join(objdir, src.replace(".cu", ".o"))
        cmd = [nvcc, *NVCC_FLAGS, *extra, "-c", os.path.join(CSRC, src), "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
            print(" ".join(cmd), flush=True)
        procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs.append(obj)
    for cmd, p in procs:
        out, _ = p.communicate()
        if verbose and out:
            print(out)
        if p.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + (out or ""))
    cmd = [nvcc, "-shared", "-o", lib, *objs, "-gencode", "arch=compute_100a,code=sm_100a"]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n" + r.stdout)
    return lib


if __name__ == "__main__":
    var = sys.argv[sys.argv.index("--variant") + 1] if "--variant" in sys.argv else ""
    defs = [a[2:] for a in sys.argv if a.startswith("-D")]
    print(build_library(force="--force" in sys.argv, verbose="-v" in sys.argv, variant=var, defines=defs))
