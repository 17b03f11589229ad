"""Latency of small batches (bs 1 / 4 / 8, Llama-3-8B heads) vs the minimum number of tiles per CTA
(fi_set_tuning min_tiles_per_cta: fewer tiles per CTA = more CTAs and more partials to merge)."""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import flashinfer_b200 as fi  # noqa: E402
from flashinfer_b200 import _lib  # noqa: E402

DEV = "cuda:0"
lib = _lib.load()
hq, hkv, d, page = 32, 8, 128, 16
for bs, kv_len in ((1, 2048), (1, 8192), (1, 32768), (4, 4096), (8, 2048), (16, 1024)):
    pages = [-(-kv_len // page)] * bs
    total = sum(pages)
    cache = torch.randn(total, 2, page, hkv, d, device=DEV, dtype=torch.bfloat16)
    indices = torch.randperm(total).to(torch.int32)
    indptr = torch.tensor([0] + list(np.cumsum(pages)), dtype=torch.int32)
    last = torch.tensor([(kv_len - 1) % page + 1] * bs, dtype=torch.int32)
    q = torch.randn(bs, hq, d, device=DEV, dtype=torch.bfloat16)
    out = torch.empty_like(q)
    row = []
    for mt in (0, 2, 4, 8):
        lib.fi_set_tuning(0, 0, 0, mt)
        w = fi.BatchDecodeWithPagedKVCacheWrapper(torch.zeros(64 << 20, dtype=torch.uint8, device=DEV), "NHD")
        w.plan(indptr, indices, last, hq, hkv, d, page, q_data_type=torch.bfloat16)
        for _ in range(5):
            w.run(q, cache, out=out)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(20):
                w.run(q, cache, out=out)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        for _ in range(10):
            g.replay()
        b.record()
        torch.cuda.synchronize()
        row.append(f"min_tiles {mt or chr(97)}: {a.elapsed_time(b) * 5:6.2f} us ({int(w._plan_info[6])} items)")
    print(f"bs {bs:2d} kv {kv_len:6d}: " + "   ".join(row))
lib.fi_set_tuning(0, 0, 0, 0)
