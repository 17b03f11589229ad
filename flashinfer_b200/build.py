"""Build libfi_b200.so (the C-ABI extension) in-tree with nvcc for sm_100a.

No JIT, no torch cpp_extension: one explicit nvcc invocation per object
(`-gencode arch=compute_100a,code=sm_100a -lineinfo`), linked into
flashinfer_b200/libfi_b200.so with the static CUDA runtime.  The reference instead
JIT-compiles per (dtype, head_dim, variant) through ninja (flashinfer/jit/core.py:250-263).

The decode kernels (114 independent template instantiations) are compiled as slices — the same
source with -DFI_SLICE_QT / -DFI_SLICE_FP8 selecting which explicit instantiations it emits
(csrc/decode_launch.h) — so a from-scratch build is bounded by the slowest slice, not by their sum.
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
# (object name, source, extra defines); the heavy slices first so the pool starts them first
UNITS = [
    ("decode_mma_f16", "decode_mma.cu", ["-DFI_SLICE_QT=0", "-DFI_SLICE_FP8=0"]),
    ("decode_mma_bf16", "decode_mma.cu", ["-DFI_SLICE_QT=1", "-DFI_SLICE_FP8=0"]),
    ("decode_mma_f16_fp8", "decode_mma.cu", ["-DFI_SLICE_QT=0", "-DFI_SLICE_FP8=1"]),
    ("decode_mma_bf16_fp8", "decode_mma.cu", ["-DFI_SLICE_QT=1", "-DFI_SLICE_FP8=1"]),
    ("decode_tc5_f16", "decode_tc5.cu", ["-DFI_SLICE_QT=0"]),
    ("decode_tc5_bf16", "decode_tc5.cu", ["-DFI_SLICE_QT=1"]),
    ("decode_variant_f16", "decode_variant.cu", ["-DFI_SLICE_QT=0"]),
    ("decode_variant_bf16", "decode_variant.cu", ["-DFI_SLICE_QT=1"]),
    ("decode", "decode.cu", []),
    ("plan", "plan.cu", []),
    ("merge", "merge.cu", []),
    ("page", "page.cu", []),
    ("gather", "gather.cu", []),
]
# the trace build is one translation unit for the decode family (decode.cu includes the slices: the
# kernels share one trace buffer)
TRACE_UNITS = [u for u in UNITS if not u[0].startswith("decode_")]
NVCC = os.environ.get("FI_B200_NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = [
    "-std=c++17", "-O3", "-lineinfo", "--expt-relaxed-constexpr",
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
    "-DFI_B200_BUILD",
]


def _digest(src: str, flags) -> str:
    """Per object: its source text, every header and slice source (decode.cu's trace build includes the
    slices), the public header, this file and the flags — editing one .cu recompiles only its objects."""
    h = hashlib.sha256()
    deps = [CSRC / src] + sorted(list(CSRC.glob("*.h")) + list(CSRC.glob("*.cuh")))
    if "-DFI_B200_TRACE" in flags and src == "decode.cu":
        deps += [CSRC / "decode_mma.cu", CSRC / "decode_tc5.cu", CSRC / "decode_variant.cu"]
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
    base = FLAGS + (["-DFI_B200_TRACE"] if trace else [])
    units = TRACE_UNITS if trace else UNITS
    flags = {name: base + defs for name, _, defs in units}
    srcs = {name: src for name, src, _ in units}
    names = [name for name, _, _ in units]
    digs = {n: _digest(srcs[n], flags[n]) for n in names}
    stamps = {n: obj_dir / (n + ".stamp") for n in names}
    objs = {n: obj_dir / (n + ".o") for n in names}

    def fresh(n: str) -> bool:
        return (not force and objs[n].exists() and stamps[n].exists()
                and stamps[n].read_text() == digs[n])

    if out.exists() and all(fresh(n) for n in names) and \
            out.stat().st_mtime >= max(objs[n].stat().st_mtime for n in names):
        return out
    if not Path(NVCC).exists():
        raise RuntimeError(f"nvcc not found at {NVCC}; cannot build {out.name}")
    obj_dir.mkdir(exist_ok=True)

    def cc(n: str) -> Path:
        if fresh(n):
            return objs[n]
        cmd = [NVCC, *flags[n], "-c", str(CSRC / srcs[n]), "-o", str(objs[n])]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {srcs[n]} ({n}):\n{r.stdout}\n{r.stderr}")
        if verbose:
            print(r.stderr)
        stamps[n].write_text(digs[n])
        return objs[n]

    with ThreadPoolExecutor(max_workers=min(len(names), os.cpu_count() or 4)) as ex:
        built = list(ex.map(cc, names))
    cmd = [NVCC, "-shared", "-o", str(out), *map(str, built), "-cudart", "static",
           "-gencode", "arch=compute_100a,code=sm_100a"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return out


if __name__ == "__main__":
    p = build(force="--force" in sys.argv, verbose="-v" in sys.argv, trace="--trace" in sys.argv)
    print(p)
