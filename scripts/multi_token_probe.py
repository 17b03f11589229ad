"""run(q_len_per_req = n) at the C2 shape (and a 70B-like head shape): time per call against n = 1.
python scripts/multi_token_probe.py"""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import flashinfer_b200 as fi

DEV = "cuda:0"


def one(hq, hkv, bs, kv_len, kv_dtype=torch.bfloat16):
    d, page = 128, 16
    pages = [-(-kv_len // page)] * bs
    total = sum(pages)
    cache = torch.randn(total, 2, page, hkv, d, device=DEV, dtype=torch.bfloat16).to(kv_dtype)
    indices = torch.randperm(total).to(torch.int32)
    indptr = torch.tensor([0] + list(np.cumsum(pages)), dtype=torch.int32)
    last = torch.full((bs,), (kv_len - 1) % page + 1, dtype=torch.int32)
    w = fi.BatchDecodeWithPagedKVCacheWrapper(torch.zeros(128 << 20, dtype=torch.uint8, device=DEV), "NHD")
    w.plan(indptr, indices, last, hq, hkv, d, page, q_data_type=torch.bfloat16, kv_data_type=kv_dtype)
    res = []
    for n in (1, 2, 4):
        q = torch.randn(bs * n, hq, d, device=DEV, dtype=torch.bfloat16)
        out = torch.empty_like(q)
        f = (lambda: w.run(q, cache, out=out)) if n == 1 else (lambda: w.run(q, cache, out=out, q_len_per_req=n))
        for _ in range(5):
            f()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        for _ in range(30):
            f()
        b.record()
        torch.cuda.synchronize()
        res.append(a.elapsed_time(b) / 30 * 1e3)
    print(f"{hq}/{hkv} heads, bs {bs}, kv {kv_len}, {str(kv_dtype)[6:]}: n=1 {res[0]:7.1f} us   n=2 {res[1]:7.1f} us   "
          f"n=4 {res[2]:7.1f} us", flush=True)


if __name__ == "__main__":
    one(32, 8, 64, 4096)
    one(64, 8, 64, 4096)
    one(32, 8, 64, 4096, torch.float8_e4m3fn)
    one(32, 8, 8, 4096)
