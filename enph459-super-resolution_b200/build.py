"""Build libsr459.so (hand-written sm_100a CUDA behind the C ABI of include/sr459.h) in-tree.

    python enph459-super-resolution_b200/build.py [--force] [-v]

nvcc cross-compiles for sm_100a without a GPU.  Every .cu file is compiled to an object of its
own (in parallel, cached under csrc/.obj by content hash), then linked; the shared library is
placed next to this file so that it travels to the GPU box with the repository snapshot.

The library carries a stamp -- the hash of every source it was built from (`sr459_source_stamp()`)
-- and `_lib.load()` compares it with the sources on disk, so a stale binary can never be tested
or benchmarked against newer sources without a signal.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(CSRC, ".obj")
LIB = os.path.join(HERE, "libsr459.so")
HEADER = os.path.join(HERE, "..", "include", "sr459.h")
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo", "-fmad=true",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
]
LINK_FLAGS = ["-shared", "-cudart", "static"]


def sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))


def _headers():
    return [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith(".h")] + [HEADER]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _digest(paths, extra=b"") -> str:
    h = hashlib.sha256(extra)
    for p in paths:
        h.update(os.path.basename(p).encode())
        with open(p, "rb") as f:
            h.update(f.read())
    return h.hexdigest()


def source_stamp() -> str:
    """Hash of everything the library is built from (sources, headers, flags): 16 hex digits."""
    paths = [os.path.join(CSRC, s) for s in sources()] + _headers()
    return _digest(paths, " ".join(NVCC_FLAGS + LINK_FLAGS).encode())[:16]


def built_stamp() -> str | None:
    """The stamp recorded beside the library by the build that produced it."""
    try:
        with open(LIB + ".stamp") as f:
            return f.read().strip()
    except OSError:
        return None


def needs_build() -> bool:
    return not os.path.exists(LIB) or built_stamp() != source_stamp()


def build(force: bool = False, verbose: bool = False) -> str:
    stamp = source_stamp()
    if not force and os.path.exists(LIB) and built_stamp() == stamp:
        return LIB
    nvcc = _nvcc()
    os.makedirs(OBJ, exist_ok=True)
    hdr_digest = _digest(_headers(), " ".join(NVCC_FLAGS).encode())
    jobs, objs = [], []
    for s in sources():
        src = os.path.join(CSRC, s)
        key = _digest([src], hdr_digest.encode())[:16]
        obj = os.path.join(OBJ, f"{s[:-3]}.{key}.o")
        objs.append(obj)
        if force or not os.path.exists(obj):
            for old in os.listdir(OBJ):
                if old.startswith(s[:-3] + ".") and old.endswith(".o"):
                    os.unlink(os.path.join(OBJ, old))
            cmd = [nvcc, *NVCC_FLAGS, f"-DSR459_SOURCE_STAMP=\"{stamp}\"" if s == "api.cu" else "-DSR459_UNUSED", "-c", src, "-o", obj]
            if verbose:
                cmd[1:1] = ["-Xptxas", "-v"]
            jobs.append((s, cmd))
    # api.cu embeds the stamp, so it is rebuilt whenever any source changed
    api_obj = [o for o in objs if os.path.basename(o).startswith("api.")][0]
    if not any(s == "api.cu" for s, _ in jobs):
        jobs.append(("api.cu", [nvcc, *NVCC_FLAGS, f"-DSR459_SOURCE_STAMP=\"{stamp}\"", "-c", os.path.join(CSRC, "api.cu"), "-o", api_obj]))

    def run(job):
        name, cmd = job
        res = subprocess.run(cmd, capture_output=True, text=True)
        return name, res

    with ThreadPoolExecutor(max_workers=min(len(jobs), os.cpu_count() or 4) or 1) as ex:
        for name, res in ex.map(run, jobs):
            if verbose:
                print(f"== {name}\n{res.stdout}{res.stderr}")
            if res.returncode != 0:
                raise RuntimeError(f"nvcc failed on {name}:\n{res.stdout}{res.stderr}")
    cmd = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", *LINK_FLAGS, "-o", LIB + ".tmp", *objs, "-ldl"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("link failed:\n" + res.stdout + res.stderr)
    os.replace(LIB + ".tmp", LIB)
    with open(LIB + ".stamp", "w") as f:
        f.write(stamp + "\n")
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
