"""Where does the fixed cost of a small decode go?  Times (events, L2 flushed, host running ahead)
an empty torch kernel, and run() on per-rank shapes (Hkv = 1) at short kv_len.  Run it under
`ncu --metrics gpu__time_duration.sum` to get the in-kernel share of each.
python scripts/overhead_probe.py [bf16|fp8]"""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import flashinfer_b200 as fi  # noqa: E402

DEV = "cuda:0"


def timed(fn, flush, n=30):
    ts = []
    for _ in range(n):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return float(np.min(ts)), float(np.median(ts))


def main():
    kind = sys.argv[1] if len(sys.argv) > 1 else "bf16"
    bs, page, d, hkv = 64, 16, 128, 1
    hq = 4 if kind == "bf16" else 8
    kv_dtype = torch.bfloat16 if kind == "bf16" else torch.float8_e4m3fn
    es = 2 if kind == "bf16" else 1
    ws = torch.zeros(128 << 20, dtype=torch.uint8, device=DEV)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)
    tiny = torch.zeros(32, device=DEV)
    print("empty torch kernel: min %.1f us median %.1f us" % timed(lambda: tiny.add_(1.0), flush))
    w = fi.BatchDecodeWithPagedKVCacheWrapper(ws, "NHD")
    for kv_len in (16, 128, 512, 1024, 4096):
        pages = [-(-kv_len // page)] * bs
        total = sum(pages)
        cache = torch.randn(total, 2, page, hkv, d, device=DEV, dtype=torch.bfloat16)
        cache = cache.to(kv_dtype) if es == 1 else cache
        indices = torch.randperm(total, device=DEV).to(torch.int32)
        indptr = torch.tensor([0] + list(np.cumsum(pages)), dtype=torch.int32)
        last = torch.tensor([(kv_len - 1) % page + 1] * bs, dtype=torch.int32)
        q = torch.randn(bs, hq, d, device=DEV, dtype=torch.bfloat16)
        w.plan(indptr, indices, last, hq, hkv, d, page, q_data_type=torch.bfloat16, kv_data_type=kv_dtype)
        for _ in range(3):
            w.run(q, cache)
        torch.cuda.synchronize()
        mn, md = timed(lambda: w.run(q, cache), flush)
        nbytes = bs * kv_len * hkv * d * 2 * es
        print(f"{kind} kv_len={kv_len:5d} {nbytes/1e6:7.2f} MB: min {mn:6.1f} us median {md:6.1f} us "
              f"(ctas {int(w._plan_info[1])}, segments {int(w._plan_info[2])}, merges {int(w._plan_info[3])})")


if __name__ == "__main__":
    main()
