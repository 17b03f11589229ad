"""NHD vs HND at the C2 (bf16) and C3 (e4m3) shapes, head-grouped order on / off.
python scripts/layout_probe.py"""
import os
import subprocess
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def one(name, layout):
    import flashinfer_b200 as fi
    DEV = "cuda:0"
    bs, kv_len, hq, hkv, fp8 = (64, 4096, 32, 8, False) if name == "c2" else (128, 8192, 64, 8, True)
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
    for _ in range(100):
        w.run(q, cache, out=out)
    b.record()
    torch.cuda.synchronize()
    us = a.elapsed_time(b) * 10
    nbytes = cache.numel() * cache.element_size()
    print(f"{name} {layout} head_group={os.environ.get('FI_B200_HEAD_GROUP', '1')}: {us:7.1f} us  {nbytes / us / 1e6:.2f} TB/s")


if __name__ == "__main__":
    if len(sys.argv) == 3:
        one(sys.argv[1], sys.argv[2])
    else:
        for name in ("c2", "c3"):
            for layout in ("NHD", "HND"):
                for hg in ("1", "0"):
                    subprocess.run([sys.executable, __file__, name, layout], env=dict(os.environ, FI_B200_HEAD_GROUP=hg))
