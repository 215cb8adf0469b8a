"""In-tree build of the native library: every ``csrc/*.cu`` / ``csrc/*.cpp`` is compiled for sm_100a and linked
into ``scaling_b200/_kernels.so`` (C ABI, loaded with ctypes — see ``scaling_b200/ops/_lib.py``).

No torch headers are involved, so a full rebuild takes seconds and works on a GPU-less box (nvcc cross-compiles).
The ``.so`` is git-ignored but travels with the working tree to the GPU box.
"""
from __future__ import annotations

import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

REPO = Path(__file__).resolve().parents[2]
CSRC = REPO / "csrc"
OUT = REPO / "scaling_b200" / "_kernels.so"
OBJ_DIR = REPO / "build" / "obj"

NVCC_FLAGS = [
    "-gencode",
    "arch=compute_100a,code=sm_100a",
    "-lineinfo",
    "-O3",
    "-std=c++17",
    "--use_fast_math",
    "-Xcompiler",
    "-fPIC",
    "-I",
    str(REPO / "include"),
    "-I",
    str(CSRC),
]
CXX_FLAGS = ["-O3", "-std=c++17", "-fPIC", "-I", str(REPO / "include"), "-I", str(CSRC)]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (Path(cand).is_file() or cand == "nvcc"):
            return cand
    raise RuntimeError("nvcc not found")


def _digest(src: Path) -> str:
    h = hashlib.sha1()
    h.update(src.read_bytes())
    for hdr in sorted(list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.h")) + list((REPO / "include").glob("*.h"))):
        h.update(hdr.read_bytes())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()[:16]


def _compile(src: Path, verbose: bool) -> Path:
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    obj = OBJ_DIR / f"{src.stem}.{_digest(src)}.o"
    if obj.is_file():
        return obj
    for old in OBJ_DIR.glob(f"{src.stem}.*.o"):
        old.unlink()
    if src.suffix == ".cu":
        cmd = [_nvcc(), *NVCC_FLAGS, "-c", str(src), "-o", str(obj)]
    else:
        cmd = ["g++", *CXX_FLAGS, "-c", str(src), "-o", str(obj)]
    if verbose:
        print(" ".join(cmd), flush=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"compile failed for {src.name}:\n{r.stdout}\n{r.stderr}")
    return obj


def build(verbose: bool = False, force: bool = False) -> Path:
    sources = sorted(CSRC.glob("*.cu")) + sorted(CSRC.glob("*.cpp"))
    if force:
        for o in OBJ_DIR.glob("*.o"):
            o.unlink()
    with ThreadPoolExecutor(max_workers=min(8, len(sources))) as ex:
        objs = list(ex.map(lambda s: _compile(s, verbose), sources))
    stamp = OBJ_DIR / "link.stamp"
    want = " ".join(o.name for o in objs)
    if OUT.is_file() and stamp.is_file() and stamp.read_text() == want and not force:
        return OUT
    # static cudart (nvcc default): the library carries the 12.9 runtime it was compiled with and shares the
    # device's primary context (and therefore streams, events and allocations) with torch's own runtime
    cmd = [_nvcc(), "-shared", "-o", str(OUT), *map(str, objs)]
    if verbose:
        print(" ".join(cmd), flush=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    stamp.write_text(want)
    return OUT


if __name__ == "__main__":
    path = build(verbose="-v" in sys.argv, force="-f" in sys.argv)
    print(path)
