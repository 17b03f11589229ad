"""Decode launch time when the L2 holds somebody else's data at launch (as inside a model, where
GEMMs run between attention calls): before every timed launch a 256 MB tensor is READ (clean
lines) or WRITTEN (dirty lines).  FI_B200_L2_HINTS=0/1 A/B.  python scripts/l2_context_probe.py c2 [shard_of]"""
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import flashinfer_b200 as fi  # noqa: E402
from flashinfer_b200 import sharding  # noqa: E402
from bench import HEAD_DIM, PAGE, page_table, workload  # noqa: E402

DEV = torch.device("cuda", 0)


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "c2"
    shard_of = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    wl = workload(name)
    sh = sharding.head_shard(wl["hq"], wl["hkv"], 0, shard_of)
    hq, hkv = sh.num_qo_heads, sh.num_kv_heads
    kv_lens = wl["kv_lens"]
    kvd = torch.float8_e4m3fn if wl["kv_dtype"] == "e4m3" else torch.bfloat16
    indptr, indices, last, total = page_table(kv_lens)
    caches = []
    for _ in range(3):
        c = torch.randn(total, 2, PAGE, hkv, HEAD_DIM, device=DEV, dtype=torch.bfloat16)
        caches.append(c.to(kvd) if kvd != torch.bfloat16 else c)
    q = torch.randn(len(kv_lens), hq, HEAD_DIM, device=DEV, dtype=torch.bfloat16)
    out = torch.empty_like(q)
    w = fi.BatchDecodeWithPagedKVCacheWrapper(torch.zeros(128 << 20, dtype=torch.uint8, device=DEV), "NHD")
    w.plan(indptr, indices, last, hq, hkv, HEAD_DIM, PAGE, q_data_type=torch.bfloat16, kv_data_type=kvd)
    other = torch.randn(64 << 20, device=DEV)  # 256 MB
    res = {}
    for mode in ("read", "write", "none"):
        ts = []
        for i in range(40):
            if mode == "read":
                other.sum()
            elif mode == "write":
                other.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            w.run(q, caches[i % 3], out=out)
            b.record()
            b.synchronize()
            ts.append(a.elapsed_time(b) * 1e3)
        res[mode] = float(np.median(ts[5:]))
    import os
    print(f"{name}/{shard_of} hints={os.environ.get('FI_B200_L2_HINTS', '0')}: " +
          "  ".join(f"after {k}: {v:7.1f} us" for k, v in res.items()))


if __name__ == "__main__":
    main()
