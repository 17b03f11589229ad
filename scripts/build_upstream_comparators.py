"""Pre-build the JIT modules of the image's upstream `flashinfer-python` (LIBRARY code, a comparator
only — never on the product path) that bench.py's `comparators` block times on the GPU box:

    fa2      BatchDecodeWithPagedKVCacheWrapper(use_tensor_cores=False)  CUDA-core decode
    fa2_tc   BatchDecodeWithPagedKVCacheWrapper(use_tensor_cores=True)   FA2 prefill kernel as decode
    trtllm-gen  the Blackwell cubin path (launcher module; cubins ship in flashinfer-cubin)

The GPU box is fresh on every call and has no network; a JIT build there would cost minutes of the
bench.  nvcc cross-compiles here, so the modules are built once in this container and copied to
baseline/_ref/fi_aot/<name>/<name>.so (git-ignored, travels with gpurun); bench.py points
flashinfer.jit.env.FLASHINFER_AOT_DIR at that directory, which makes JitSpec.build_and_load() load
the file directly (flashinfer/jit/core.py `is_aot`).

    FLASHINFER_CUDA_ARCH_LIST=10.0a python scripts/build_upstream_comparators.py
"""
from __future__ import annotations

import os
import shutil
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
OUT = ROOT / "baseline" / "_ref" / "fi_aot"
os.environ.setdefault("FLASHINFER_CUDA_ARCH_LIST", "10.0a")
os.environ.setdefault("FLASHINFER_WORKSPACE_BASE", "/tmp/fi_b200_upstream_ws")
os.environ.setdefault("MAX_JOBS", str(os.cpu_count() or 4))


def main() -> None:
    import torch
    from flashinfer.jit.attention import modules as M

    bf, f16, e4 = torch.bfloat16, torch.float16, torch.float8_e4m3fn
    i32 = torch.int32
    specs = []
    for q, kv in ((bf, bf), (bf, e4), (f16, f16)):
        # (q, kv, o, idtype, D_qk, D_vo, pos_encoding NONE, sliding window, soft cap)
        specs.append(M.gen_batch_decode_module(q, kv, q, i32, 128, 128, 0, False, False))
    for q, kv in ((bf, bf), (bf, e4), (f16, f16)):
        specs.append(M.gen_batch_prefill_module("fa2", q, kv, q, i32, 128, 128, 0, False, False, False))
    specs.append(M.gen_trtllm_gen_fmha_module())
    only = set(sys.argv[1:])
    OUT.mkdir(parents=True, exist_ok=True)
    for s in specs:
        if only and not any(o in s.name for o in only):
            continue
        dst = OUT / s.name / f"{s.name}.so"
        if dst.exists():
            print(f"have   {s.name}")
            continue
        t0 = time.time()
        s.build(verbose=False)
        dst.parent.mkdir(parents=True, exist_ok=True)
        shutil.copy2(s.jit_library_path, dst)
        print(f"built  {s.name}  {time.time() - t0:.0f} s  {dst.stat().st_size >> 10} KiB", flush=True)


if __name__ == "__main__":
    main()
