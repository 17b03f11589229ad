"""Fixed-overhead vs streaming-rate fit: time run() for per-rank shapes (Hkv = 1) at several
kv_len and fit t = a + bytes / rate.  python scripts/size_sweep.py [bf16|fp8] [bs]"""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import flashinfer_b200 as fi  # noqa: E402

DEV = "cuda:0"


def main():
    kind = sys.argv[1] if len(sys.argv) > 1 else "bf16"
    bs = int(sys.argv[2]) if len(sys.argv) > 2 else 64
    page, d, hkv = 16, 128, 1
    hq = 4 if kind == "bf16" else 8
    kv_dtype = torch.bfloat16 if kind == "bf16" else torch.float8_e4m3fn
    es = 2 if kind == "bf16" else 1
    ws = torch.zeros(128 << 20, dtype=torch.uint8, device=DEV)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)
    w = fi.BatchDecodeWithPagedKVCacheWrapper(ws, "NHD")
    xs, ys = [], []
    for kv_len in (1024, 2048, 4096, 8192, 16384, 32768):
        kv_lens = [kv_len] * bs
        pages = [-(-l // page) for l in kv_lens]
        total = sum(pages)
        cache = torch.randn(total, 2, page, hkv, d, device=DEV, dtype=torch.bfloat16)
        cache = cache.to(kv_dtype) if es == 1 else cache
        indices = torch.randperm(total, device=DEV).to(torch.int32)
        indptr = torch.tensor([0] + list(np.cumsum(pages)), dtype=torch.int32)
        last = torch.tensor([((l - 1) % page + 1) for l in kv_lens], dtype=torch.int32)
        q = torch.randn(bs, hq, d, device=DEV, dtype=torch.bfloat16)
        w.plan(indptr, indices, last, hq, hkv, d, page, q_data_type=torch.bfloat16, kv_data_type=kv_dtype)
        for _ in range(3):
            w.run(q, cache)
        torch.cuda.synchronize()
        ts = []
        for _ in range(30):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            w.run(q, cache)
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3)
        nbytes = sum(kv_lens) * hkv * d * 2 * es + bs * hq * d * 4
        t = float(np.min(ts))
        xs.append(nbytes)
        ys.append(t)
        print(f"{kind} bs={bs} kv_len={kv_len:6d}: {nbytes/1e6:8.1f} MB  min {t:7.1f} us  median {np.median(ts):7.1f} us "
              f"-> {nbytes/t/1e6:.2f} TB/s  (segments {int(w._plan_info[2])}, merges {int(w._plan_info[3])})")
        del cache
    b, a = np.polyfit(np.array(xs), np.array(ys), 1)
    print(f"fit: t = {a:.1f} us + bytes / {1/b/1e6:.2f} TB/s")


if __name__ == "__main__":
    main()
