"""In-tree build of libb200_transducer.so (nvcc, sm_100a only) and of the C oracle.

Run as ``python -m wenet_celoss_b200.build`` or through ``__graft_entry__.build()``.
The .so lands next to this file so it travels with the repo snapshot to the GPU box.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from pathlib import Path

PKG = Path(__file__).resolve().parent
ROOT = PKG.parent
CSRC = PKG / "csrc"
LIB = PKG / "libb200_transducer.so"
OBJ_DIR = PKG / "_obj"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC",
    "-DB200_ARCH=100",
    "--expt-relaxed-constexpr",
]


def _nvcc() -> str:
    nv = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nv):
        raise RuntimeError("nvcc not found: libb200_transducer.so cannot be built")
    return nv


def _digest(paths) -> str:
    h = hashlib.sha256()
    h.update(" ".join(NVCC_FLAGS).encode())
    for p in sorted(paths):
        h.update(p.name.encode())
        h.update(p.read_bytes())
    return h.hexdigest()


def build_library(force: bool = False, verbose: bool = False) -> Path:
    """Compile every csrc/*.cu for sm_100a and link the shared library."""
    sources = sorted(CSRC.glob("*.cu"))
    headers = (sorted(CSRC.glob("*.cuh")) + sorted(CSRC.glob("*.inc")) +
               sorted((ROOT / "include").glob("*.h")))
    OBJ_DIR.mkdir(exist_ok=True)
    hdr_digest = _digest(headers)
    nvcc = _nvcc()
    objs = []
    procs = []
    for src in sources:
        obj = OBJ_DIR / (src.stem + ".o")
        stamp = OBJ_DIR / (src.stem + ".stamp")
        want = _digest([src]) + hdr_digest
        objs.append(obj)
        if not force and obj.exists() and stamp.exists() and stamp.read_text() == want:
            continue
        cmd = [nvcc, *NVCC_FLAGS, "-c", str(src), "-o", str(obj)]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
            print(" ".join(cmd), flush=True)
        procs.append((src, stamp, want, subprocess.Popen(cmd, stdout=subprocess.PIPE,
                                                          stderr=subprocess.STDOUT, text=True)))
    relink = force or not LIB.exists() or bool(procs)
    for src, stamp, want, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src.name}:\n{out}")
        if verbose and out:
            print(out)
        stamp.write_text(want)
    if relink:
        cmd = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", str(LIB),
               *map(str, objs), "-lcudart"]
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}")
    return LIB


def build_oracle(force: bool = False) -> Path:
    """Compile oracle/'s C restatement (test infrastructure, not the product)."""
    odir = ROOT / "oracle"
    src = odir / "lattice_ref.c"
    out = odir / "_build" / "liboracle_ref.so"
    if not src.exists():
        return out
    out.parent.mkdir(exist_ok=True)
    if force or not out.exists() or out.stat().st_mtime < src.stat().st_mtime:
        cmd = ["gcc", "-O2", "-fPIC", "-shared", "-fopenmp", "-o", str(out), str(src), "-lm"]
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"gcc failed on oracle:\n{r.stdout}")
    return out


if __name__ == "__main__":
    force = "--force" in sys.argv
    verbose = "-v" in sys.argv
    print(build_library(force=force, verbose=verbose))
    print(build_oracle(force=force))
