"""Event time of run() for one shape (L2 flushed between launches, host running ahead).
python scripts/shape_time.py bf16|fp8 kv_len hkv [bs] — scheduler knobs come from the environment
(FI_B200_SCHED_F0 / _G / _TAIL / _LOOKAHEAD), so A/B runs are separate processes."""
import os
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import flashinfer_b200 as fi  # noqa: E402

DEV = "cuda:0"


def main():
    kind, kv_len, hkv = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
    bs = int(sys.argv[4]) if len(sys.argv) > 4 else 64
    pos_mode = sys.argv[5] if len(sys.argv) > 5 else "NONE"
    page, d = 16, 128
    hq = hkv * (4 if kind == "bf16" else 8)
    kv_dtype = torch.bfloat16 if kind == "bf16" else torch.float8_e4m3fn
    es = 2 if kind == "bf16" else 1
    ws = torch.zeros(128 << 20, dtype=torch.uint8, device=DEV)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)
    w = fi.BatchDecodeWithPagedKVCacheWrapper(ws, "NHD")
    pages = [-(-kv_len // page)] * bs
    total = sum(pages)
    cache = torch.randn(total, 2, page, hkv, d, device=DEV, dtype=torch.bfloat16)
    cache = cache.to(kv_dtype) if es == 1 else cache
    indices = torch.randperm(total, device=DEV).to(torch.int32)
    indptr = torch.tensor([0] + list(np.cumsum(pages)), dtype=torch.int32)
    last = torch.tensor([(kv_len - 1) % page + 1] * bs, dtype=torch.int32)
    q = torch.randn(bs, hq, d, device=DEV, dtype=torch.bfloat16)
    w.plan(indptr, indices, last, hq, hkv, d, page, q_data_type=torch.bfloat16, kv_data_type=kv_dtype,
           pos_encoding_mode=pos_mode)
    for _ in range(3):
        w.run(q, cache)
    torch.cuda.synchronize()
    ts = []
    for _ in range(40):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        w.run(q, cache)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    nbytes = bs * kv_len * hkv * d * 2 * es
    knobs = " ".join(f"{k[14:]}={v}" for k, v in sorted(os.environ.items()) if k.startswith("FI_B200_SCHED_"))
    print(f"{kind} kv={kv_len} hkv={hkv} bs={bs} {pos_mode} [{knobs}]: min {min(ts):7.1f} median {np.median(ts):7.1f} us "
          f"-> {nbytes / np.median(ts) / 1e6:.2f} TB/s (items {int(w._plan_info[6])}, segs {int(w._plan_info[2])}, "
          f"merges {int(w._plan_info[3])})")


if __name__ == "__main__":
    main()
