"""append_paged_kv_cache throughput: prefill-sized (16 requests x 4096 new tokens) and decode-sized
(256 requests x 1 token) appends, Llama-3-8B KV shape (8 heads x 128, bf16, page 16, NHD)."""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import flashinfer_b200 as fi  # noqa: E402

DEV = "cuda:0"
hkv, d, page = 8, 128, 16
for bs, new, old in ((16, 4096, 0), (64, 512, 2048), (256, 1, 4095)):
    lens = [old + new] * bs
    pages = [-(-l // page) for l in lens]
    total = sum(pages)
    cache = torch.zeros(total, 2, page, hkv, d, device=DEV, dtype=torch.bfloat16)
    kv_indices = torch.randperm(total, device=DEV).to(torch.int32)
    kv_indptr = torch.tensor([0] + list(np.cumsum(pages)), dtype=torch.int32, device=DEV)
    kv_last = torch.tensor([(l - 1) % page + 1 for l in lens], dtype=torch.int32, device=DEV)
    nnz = bs * new
    k = torch.randn(nnz, hkv, d, device=DEV, dtype=torch.bfloat16)
    v = torch.randn(nnz, hkv, d, device=DEV, dtype=torch.bfloat16)
    append_indptr = torch.arange(bs + 1, device=DEV, dtype=torch.int32) * new
    seq_lens = fi.get_seq_lens(kv_indptr, kv_last, page)
    bi, pos = fi.get_batch_indices_positions(append_indptr, seq_lens, nnz)

    def call():
        fi.append_paged_kv_cache(k, v, bi, pos, cache, kv_indices, kv_indptr, kv_last)

    for _ in range(5):
        call()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    n = 50
    a.record()
    for _ in range(n):
        call()
    b.record()
    torch.cuda.synchronize()
    us = a.elapsed_time(b) * 1e3 / n
    nbytes = 2 * 2 * nnz * hkv * d * 2  # K and V, read + written
    print(f"bs {bs:3d} x {new:4d} new tokens (nnz {nnz:6d}): {us:8.1f} us  {nbytes / us / 1e6:.2f} TB/s (read + write)")
