"""Timing of the cascade wrapper (one decode request per query row, re-reads served by the L2) against
the same attention as ONE flat batch decode over per-request copies of the page lists.
Shared prefix of P tokens for all B requests + a unique suffix of U tokens each, Llama-3-8B heads.
python scripts/cascade_probe.py"""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import flashinfer_b200 as fi

DEV = "cuda:0"
hq, hkv, d, page = 32, 8, 128, 16


def timed(fn, n=30):
    for _ in range(5):
        fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3


def one(B, P, U):
    pp, up = P // page, -(-U // page)
    total = pp + B * up
    cache = torch.randn(total, 2, page, hkv, d, device=DEV, dtype=torch.bfloat16)
    q = torch.randn(B, hq, d, device=DEV, dtype=torch.bfloat16)
    ws = torch.zeros(128 << 20, dtype=torch.uint8, device=DEV)
    shared = torch.arange(pp, dtype=torch.int32)
    uniq = [torch.arange(pp + r * up, pp + (r + 1) * up, dtype=torch.int32) for r in range(B)]
    last_u = (U - 1) % page + 1
    # cascade: level 0 = the shared prefix (one group, B rows), level 1 = the unique suffixes
    w = fi.MultiLevelCascadeAttentionWrapper(2, ws, "NHD")
    qo = [torch.tensor([0, B], dtype=torch.int32), torch.arange(B + 1, dtype=torch.int32)]
    kvp = [torch.tensor([0, pp], dtype=torch.int32), torch.arange(B + 1, dtype=torch.int32) * up]
    kvi = [shared, torch.cat(uniq)]
    lpl = [torch.tensor([page], dtype=torch.int32), torch.full((B,), last_u, dtype=torch.int32)]
    w.plan(qo, kvp, kvi, lpl, hq, hkv, d, page, q_data_type=torch.bfloat16)
    t_cascade = timed(lambda: w.run(q, cache))
    # flat: every request lists the shared pages and its own
    f = fi.BatchDecodeWithPagedKVCacheWrapper(ws, "NHD")
    indptr = torch.arange(B + 1, dtype=torch.int32) * (pp + up)
    indices = torch.cat([torch.cat([shared, u]) for u in uniq])
    f.plan(indptr, indices, torch.full((B,), last_u, dtype=torch.int32), hq, hkv, d, page, q_data_type=torch.bfloat16)
    t_flat = timed(lambda: f.run(q, cache))
    o1, o2 = w.run(q, cache), f.run(q, cache)
    err = (o1.float() - o2.float()).abs().max().item()
    shared_mb = pp * 2 * page * hkv * d * 2 / 1e6
    uniq_mb = B * up * 2 * page * hkv * d * 2 / 1e6
    print(f"B={B:4d} shared {P:6d} tok ({shared_mb:6.1f} MB) + unique {U:5d} tok ({uniq_mb:7.1f} MB total): cascade "
          f"{t_cascade:7.1f} us, flat decode {t_flat:7.1f} us, max|do| {err:.1e}", flush=True)


if __name__ == "__main__":
    for B, P, U in ((64, 8192, 256), (64, 32768, 256), (16, 8192, 1024), (256, 2048, 512), (64, 8192, 4096)):
        one(B, P, U)
