"""Launches every kernel of libfi_b200.so that is not the main decode kernel once at a bandwidth-relevant
size (for scripts/ncu_round2.sh), and prints achieved GB/s against algorithmic bytes (inputs + outputs)
when run without a profiler: decode_merge_kernel (unfused split-KV merge and single decode), the public
merge operators, append_paged_kv_cache, get_batch_indices_positions, the CUDA-core RoPE variant kernel
(fp8 KV + ROPE_LLAMA), head_dim 64 / 256 instantiations."""
import os
import sys
import time
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import flashinfer_b200 as fi  # noqa: E402

DEV = "cuda:0"


def timed(fn, n=20):
    if os.environ.get("FI_PROBE_ONCE") == "1":  # under ncu: one launch of every kernel is enough
        fn()
        torch.cuda.synchronize()
        return float("nan")
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3


def main():
    torch.manual_seed(0)
    bf = torch.bfloat16
    # ---- public merge operators: seq x heads x D rows
    seq, H, D = 8192, 32, 128
    va, vb = torch.randn(seq, H, D, device=DEV, dtype=bf), torch.randn(seq, H, D, device=DEV, dtype=bf)
    sa, sb = torch.randn(seq, H, device=DEV), torch.randn(seq, H, device=DEV)
    us = timed(lambda: fi.merge_state(va, sa, vb, sb))
    byt = 3 * va.numel() * 2 + 3 * sa.numel() * 4
    print(f"merge_state            {seq}x{H}x{D}: {us:7.1f} us  {byt / us / 1e3:7.1f} GB/s")
    v2 = va.clone()
    us = timed(lambda: fi.merge_state_in_place(v2, sa.clone(), vb, sb))
    print(f"merge_state_in_place   {seq}x{H}x{D}: {us:7.1f} us  (includes an lse clone)")
    vs = torch.randn(seq, 4, H, D, device=DEV, dtype=bf)
    ss = torch.randn(seq, 4, H, device=DEV)
    us = timed(lambda: fi.merge_states(vs, ss))
    byt = vs.numel() * 2 + ss.numel() * 4 + seq * H * D * 2 + seq * H * 4
    print(f"merge_states (n = 4)   {seq}x{H}x{D}: {us:7.1f} us  {byt / us / 1e3:7.1f} GB/s")
    # ---- append_paged_kv_cache + get_batch_indices_positions (16 requests x 4096 new tokens)
    bs, new, page, hkv = 16, 4096, 16, 8
    pages = bs * new // page
    cache = torch.zeros(pages, 2, page, hkv, D, device=DEV, dtype=bf)
    k_new = torch.randn(bs * new, hkv, D, device=DEV, dtype=bf)
    v_new = torch.randn_like(k_new)
    indptr = (torch.arange(bs + 1, dtype=torch.int32, device=DEV) * (new // page))
    indices = torch.randperm(pages, device=DEV).to(torch.int32)
    last = torch.full((bs,), page, dtype=torch.int32, device=DEV)
    append_indptr = (torch.arange(bs + 1, dtype=torch.int32, device=DEV) * new)
    seq_lens = fi.get_seq_lens(indptr, last, page)
    us = timed(lambda: fi.get_batch_indices_positions(append_indptr, seq_lens, bs * new))
    print(f"get_batch_indices_positions {bs * new} tokens: {us:7.1f} us")
    bi, pos = fi.get_batch_indices_positions(append_indptr, seq_lens, bs * new)
    us = timed(lambda: fi.append_paged_kv_cache(k_new, v_new, bi, pos, cache, indices, indptr, last))
    byt = 2 * (k_new.numel() * 2) * 2
    print(f"append_paged_kv_cache  {bs}x{new} tokens: {us:7.1f} us  {byt / us / 1e3:7.1f} GB/s (read + write)")
    # ---- single decode (implicit plan + decode_merge_kernel), head dims 64 / 128 / 256
    for d in (64, 128, 256):
        q = torch.randn(32, d, device=DEV, dtype=bf)
        k = torch.randn(32768, 8, d, device=DEV, dtype=bf)
        v = torch.randn_like(k)
        us = timed(lambda: fi.single_decode_with_kv_cache(q, k, v))
        print(f"single_decode D={d:3d} kv 32768: {us:7.1f} us  {2 * k.numel() * 2 / us / 1e3:7.1f} GB/s")
    # ---- batch decode with the split-KV merge as a separate launch (A/B switch) needs a fresh process;
    # the CUDA-core RoPE variant (fp8 KV + ROPE_LLAMA) at a C3-rank-like shape
    bs, kv_len, hq, hkv = 32, 4096, 8, 1
    npg = bs * kv_len // page
    c8 = torch.randn(npg, 2, page, hkv, D, device=DEV, dtype=bf).to(torch.float8_e4m3fn)
    ip = torch.arange(bs + 1, dtype=torch.int32) * (kv_len // page)
    ix = torch.randperm(npg).to(torch.int32)
    lp = torch.full((bs,), page, dtype=torch.int32)
    w = fi.BatchDecodeWithPagedKVCacheWrapper(torch.zeros(64 << 20, dtype=torch.uint8, device=DEV), "NHD")
    w.plan(ip, ix, lp, hq, hkv, D, page, pos_encoding_mode="ROPE_LLAMA", q_data_type=bf,
           kv_data_type=torch.float8_e4m3fn)
    q8 = torch.randn(bs, hq, D, device=DEV, dtype=bf)
    us = timed(lambda: w.run(q8, c8))
    print(f"fp8 KV + ROPE_LLAMA (CUDA-core variant) bs{bs} kv{kv_len}: {us:7.1f} us  {c8.numel() / us / 1e3:7.1f} GB/s")


if __name__ == "__main__":
    main()
