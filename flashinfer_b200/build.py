"""Build libfi_b200.so (the C-ABI extension) in-tree with nvcc for sm_100a.

No JIT, no torch cpp_extension: one explicit nvcc invocation per translation unit
(`-gencode arch=compute_100a,code=sm_100a -lineinfo`), linked into
flashinfer_b200/libfi_b200.so with the static CUDA runtime.  The reference instead
JIT-compiles per (dtype, head_dim, variant) through ninja (flashinfer/jit/core.py:250-263).
"""
from __future__ import annotations

import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

HERE = Path(__file__).resolve().parent
CSRC = HERE / "csrc"
OUT = HERE / "libfi_b200.so"
OBJ = HERE / "build"
SOURCES = ["plan.cu", "decode.cu", "merge.cu", "page.cu", "gather.cu"]
NVCC = os.environ.get("FI_B200_NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = [
    "-std=c++17", "-O3", "-lineinfo", "--expt-relaxed-constexpr",
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
    "-DFI_B200_BUILD",
]


def _digest(src: str, flags) -> str:
    """Per translation unit: its own text, every header (any TU may include any of them), the
    public header, this file and the flags — editing one .cu recompiles only that .cu."""
    h = hashlib.sha256()
    deps = [CSRC / src] + sorted(list(CSRC.glob("*.h")) + list(CSRC.glob("*.cuh")))
    for p in deps + [HERE.parent / "include" / "fi_b200.h", Path(__file__)]:
        h.update(p.name.encode())
        h.update(p.read_bytes())
    h.update(" ".join(flags).encode())
    return h.hexdigest()


def build(force: bool = False, verbose: bool = False, trace: bool = False) -> Path:
    """trace=True builds libfi_b200_trace.so (-DFI_B200_TRACE: per-CTA phase timestamps for
    scripts/trace_probe.py); the shipped library never carries the trace code."""
    out = HERE / "libfi_b200_trace.so" if trace else OUT
    obj_dir = HERE / "build_trace" if trace else OBJ
    flags = FLAGS + (["-DFI_B200_TRACE"] if trace else [])
    digs = {src: _digest(src, flags) for src in SOURCES}
    stamps = {src: obj_dir / (src + ".stamp") for src in SOURCES}

    def fresh(src: str) -> bool:
        return (not force and (obj_dir / (src + ".o")).exists() and stamps[src].exists()
                and stamps[src].read_text() == digs[src])

    if out.exists() and all(fresh(src) for src in SOURCES) and \
            out.stat().st_mtime >= max((obj_dir / (src + ".o")).stat().st_mtime for src in SOURCES):
        return out
    if not Path(NVCC).exists():
        raise RuntimeError(f"nvcc not found at {NVCC}; cannot build {out.name}")
    obj_dir.mkdir(exist_ok=True)

    def cc(src: str) -> Path:
        obj = obj_dir / (src + ".o")
        if fresh(src):
            return obj
        cmd = [NVCC, *flags, "-c", str(CSRC / src), "-o", str(obj)]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            print(r.stderr)
        stamps[src].write_text(digs[src])
        return obj

    with ThreadPoolExecutor(max_workers=len(SOURCES)) as ex:
        objs = list(ex.map(cc, SOURCES))
    cmd = [NVCC, "-shared", "-o", str(out), *map(str, objs), "-cudart", "static",
           "-gencode", "arch=compute_100a,code=sm_100a"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return out


if __name__ == "__main__":
    p = build(force="--force" in sys.argv, verbose="-v" in sys.argv, trace="--trace" in sys.argv)
    print(p)
