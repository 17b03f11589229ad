"""Run one config a few times (for ncu): python scripts/one_shot.py c2r [NHD] [iters]"""
import sys
from pathlib import Path
import numpy as np
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
sys.path.insert(0, str(Path(__file__).resolve().parent))
import flashinfer_b200 as fi
from perf_probe import config
name = sys.argv[1]; layout = sys.argv[2] if len(sys.argv) > 2 else "NHD"; iters = int(sys.argv[3]) if len(sys.argv) > 3 else 6
c = config(name); page, d = 16, 128
kv_lens, hq, hkv, kv_dtype = c["kv_lens"], c["hq"], c["hkv"], c["kv_dtype"]
pages = [-(-l // page) for l in kv_lens]; total = sum(pages)
shape = (total, 2, page, hkv, d) if layout == "NHD" else (total, 2, hkv, page, d)
cache = torch.randn(shape, device="cuda", dtype=torch.bfloat16)
if kv_dtype == torch.float8_e4m3fn: cache = cache.to(kv_dtype)
indices = torch.randperm(total, device="cuda").to(torch.int32)
indptr = torch.tensor([0] + list(np.cumsum(pages)), dtype=torch.int32)
last = torch.tensor([((l - 1) % page + 1) for l in kv_lens], dtype=torch.int32)
q = torch.randn(len(kv_lens), hq, d, device="cuda", dtype=torch.bfloat16)
ws = torch.zeros(128 << 20, dtype=torch.uint8, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
w = fi.BatchDecodeWithPagedKVCacheWrapper(ws, layout)
w.plan(indptr, indices, last, hq, hkv, d, page, q_data_type=torch.bfloat16, kv_data_type=kv_dtype)
for _ in range(iters):
    flush.zero_()
    w.run(q, cache)
torch.cuda.synchronize()
print("ok")
