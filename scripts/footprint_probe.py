"""Does throughput depend on the size of the cache?  C2-like (bf16) and C3-like (e4m3) head shapes
at 1, 2 and 4 GB of KV, back-to-back launches.  python scripts/footprint_probe.py"""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import flashinfer_b200 as fi

DEV = "cuda:0"


def one(fp8, bs, kv_len, layout="NHD"):
    hq, hkv = (64, 8) if fp8 else (32, 8)
    d, page = 128, 16
    pages = [kv_len // page] * bs
    total = sum(pages)
    shape = (total, 2, page, hkv, d) if layout == "NHD" else (total, 2, hkv, page, d)
    cache = torch.randn(*shape, device=DEV, dtype=torch.bfloat16)
    if fp8:
        cache = cache.to(torch.float8_e4m3fn)
    indices = torch.randperm(total).to(torch.int32)
    indptr = torch.tensor([0] + list(np.cumsum(pages)), dtype=torch.int32)
    last = torch.full((bs,), page, dtype=torch.int32)
    q = torch.randn(bs, hq, d, device=DEV, dtype=torch.bfloat16)
    out = torch.empty_like(q)
    w = fi.BatchDecodeWithPagedKVCacheWrapper(torch.zeros(128 << 20, dtype=torch.uint8, device=DEV), layout)
    w.plan(indptr, indices, last, hq, hkv, d, page, q_data_type=torch.bfloat16, kv_data_type=cache.dtype)
    for _ in range(5):
        w.run(q, cache, out=out)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    a.record()
    for _ in range(50):
        w.run(q, cache, out=out)
    b.record()
    torch.cuda.synchronize()
    us = a.elapsed_time(b) * 20
    nbytes = cache.numel() * cache.element_size()
    print(f"{'e4m3' if fp8 else 'bf16'} {layout} bs={bs:4d} kv={kv_len:6d}: {nbytes / 2**30:5.2f} GiB {us:8.1f} us  "
          f"{nbytes / us / 1e6:.2f} TB/s", flush=True)
    del cache


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "fp8":  # fp8 only, more shapes (planner A/B runs)
        for layout in ("NHD", "HND"):
            for bs, kv in ((32, 8192), (64, 8192), (128, 8192), (192, 8192), (256, 8192), (128, 2048), (128, 4096),
                           (16, 65536), (37, 20000)):
                one(True, bs, kv, layout)
        sys.exit(0)
    for fp8, kv in ((False, 4096), (True, 8192)):
        for bs in (32, 64, 128, 256, 512):
            one(fp8, bs, kv)
    one(True, 64, 8192, "HND")
    one(True, 256, 8192, "HND")
