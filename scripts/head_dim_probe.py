"""Streaming rate of the 16-bit decode kernel at head_dim 64 / 128 / 256 (1 GiB of KV each, Hkv = 8, G = 4).
python scripts/head_dim_probe.py [fp8]"""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import flashinfer_b200 as fi  # noqa: E402

DEV = "cuda:0"
fp8 = len(sys.argv) > 1 and sys.argv[1] == "fp8"
for d in ((128, 256) if fp8 else (64, 128, 256)):
    hkv, hq, page, bs = 8, 32, 16, 64
    es = 1 if fp8 else 2
    kv_len = (1 << 30) // (bs * hkv * d * 2 * es)
    pages = [-(-kv_len // page)] * bs
    total = sum(pages)
    cache = torch.randn(total, 2, page, hkv, d, device=DEV, dtype=torch.bfloat16)
    if fp8:
        cache = cache.to(torch.float8_e4m3fn)
    indices = torch.randperm(total).to(torch.int32)
    indptr = torch.tensor([0] + list(np.cumsum(pages)), dtype=torch.int32)
    last = torch.tensor([(kv_len - 1) % page + 1] * bs, dtype=torch.int32)
    q = torch.randn(bs, hq, d, device=DEV, dtype=torch.bfloat16)
    out = torch.empty_like(q)
    w = fi.BatchDecodeWithPagedKVCacheWrapper(torch.zeros(128 << 20, dtype=torch.uint8, device=DEV), "NHD")
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
    nbytes = bs * kv_len * hkv * d * 2 * es
    print(f"{'e4m3' if fp8 else 'bf16'} D={d} kv_len={kv_len}: {us:7.1f} us  {nbytes / us / 1e6:.2f} TB/s  "
          f"(ctas {int(w._plan_info[1])}, stages {int(w._plan_info[22])})")
    del cache
