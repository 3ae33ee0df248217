"""Builds the in-tree CUDA library for sm_100a:  python -m lmc_atomi_b200.build"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
# LMC_B200_LIB / LMC_NVCC_EXTRA: build (and load, see _lib.py) an experimental variant next to the default library
OUT = os.environ.get("LMC_B200_LIB", os.path.join(HERE, "_lmc_b200.so"))
EXTRA = os.environ.get("LMC_NVCC_EXTRA", "").split()
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-diag-suppress", "177"]


def sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))


def _includes(path, seen):
    """Files `path` includes with #include "..." (recursively): the per-object dependency list."""
    import re
    if path in seen or not os.path.isfile(path):
        return seen
    seen.add(path)
    for m in re.finditer(r'^\s*#\s*include\s+"([^"]+)"', open(path).read(), flags=re.M):
        _includes(os.path.normpath(os.path.join(os.path.dirname(path), m.group(1))), seen)
    return seen


def build(force=False, verbose=False):
    srcs = sources()
    nvcc = os.environ.get("NVCC", "nvcc")
    objs = []
    procs = []
    bdir = os.path.join(HERE, "build", os.path.basename(OUT)[:-3])
    os.makedirs(bdir, exist_ok=True)
    flags_tag = os.path.join(bdir, "flags.txt")
    flags = " ".join([*NVCC_FLAGS, *EXTRA])
    if not os.path.isfile(flags_tag) or open(flags_tag).read() != flags:
        force = True
    for s in srcs:
        o = os.path.join(bdir, os.path.basename(s)[:-3] + ".o")
        objs.append(o)
        deps = _includes(s, set())
        if not force and os.path.isfile(o) and all(os.path.getmtime(o) >= os.path.getmtime(d) for d in deps):
            continue      # object is newer than its source and every header it includes
        cmd = [nvcc, *NVCC_FLAGS, *EXTRA, "-c", s, "-o", o] + (["-Xptxas", "-v"] if verbose else [])
        procs.append((cmd, subprocess.Popen(cmd)))
    if not procs and os.path.isfile(OUT) and all(os.path.getmtime(OUT) >= os.path.getmtime(o) for o in objs):
        return OUT
    for cmd, p in procs:
        if p.wait() != 0:
            raise RuntimeError("nvcc failed: " + " ".join(cmd))
    cmd = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", OUT, *objs, "-lcudart"]
    subprocess.check_call(cmd)
    with open(flags_tag, "w") as f:
        f.write(flags)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
