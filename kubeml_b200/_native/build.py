"""Build the native runtime in-tree with plain nvcc for sm_100a.

    python -m kubeml_b200._native.build [--force] [--verbose]

Every translation unit is compiled with `-gencode arch=compute_100a,code=sm_100a -lineinfo`
(tcgen05 / TMEM / TMA need the arch-specific target; -lineinfo lets `ncu --import-source on` map
SASS back to source) and linked, with the static CUDA runtime, into
kubeml_b200/_native/libkubeml_b200.so.  Driver-API symbols are resolved lazily at run time
(csrc/driver_api.cpp) so the library also loads on the GPU-less build machine.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
CSRC = os.path.join(ROOT, "csrc")
INCLUDE = os.path.join(ROOT, "include")
OUT = os.path.join(HERE, "libkubeml_b200.so")
OBJ_DIR = os.path.join(ROOT, "build", "obj")

ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=default",
          "--expt-relaxed-constexpr", f"-I{INCLUDE}", f"-I{CSRC}"]


def nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.isfile(cand):
            return cand
    raise RuntimeError("nvcc not found")


def sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cpp")))


def _digest(paths) -> str:
    h = hashlib.sha256()
    for p in sorted(paths):
        with open(p, "rb") as f:
            h.update(p.encode())
            h.update(f.read())
    return h.hexdigest()


def _deps():
    hdrs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    hdrs += [os.path.join(INCLUDE, f) for f in os.listdir(INCLUDE) if f.endswith(".h")]
    return hdrs


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJ_DIR, exist_ok=True)
    srcs = sources()
    hdr_digest = _digest(_deps())
    cc = nvcc()
    objs, jobs = [], []
    for s in srcs:
        o = os.path.join(OBJ_DIR, os.path.basename(s) + ".o")
        stamp = o + ".sha"
        want = _digest([s]) + hdr_digest
        objs.append(o)
        have = open(stamp).read() if os.path.isfile(stamp) else ""
        if force or have != want or not os.path.isfile(o):
            cmd = [cc, *ARCH_FLAGS, *COMMON, "-x", "cu", "-c", s, "-o", o]
            jobs.append((cmd, stamp, want))

    def run(job):
        cmd, stamp, want = job
        if verbose:
            print(" ".join(cmd), flush=True)
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed:\n{' '.join(cmd)}\n{r.stdout}\n{r.stderr}")
        with open(stamp, "w") as f:
            f.write(want)

    if jobs:
        with ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            list(ex.map(run, jobs))
    if jobs or force or not os.path.isfile(OUT):
        link = [cc, *ARCH_FLAGS, "-shared", "-o", OUT, *objs, "-cudart", "static", "-ldl", "-lpthread"]
        if verbose:
            print(" ".join(link), flush=True)
        r = subprocess.run(link, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return OUT


if __name__ == "__main__":
    path = build(force="--force" in sys.argv, verbose="--verbose" in sys.argv)
    print(path)
