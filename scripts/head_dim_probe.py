"""Streaming rate of the 16-bit decode kernel at head_dim 64 / 128 / 256 (1 GiB of KV each, Hkv = 8, G = 4).
python scripts/head_dim_probe.py [fp8 [head_dim]]"""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import flashinfer_b200 as fi  # noqa: E402

DEV = "cuda:0"
fp8 = len(sys.argv) > 1 and sys.argv[1] == "fp8"
only = int(sys.argv[2]) if len(sys.argv) > 2 else 0  # e.g. `fp8 256`: that head_dim only (ncu captures)
for d in ((only,) if only else ((128, 256) if fp8 else (64, 128, 256))):
    hkv, hq, page, bs = 8, 32, 16, 64
    es = 1 if fp8 else 2
    kv_len = (1 << 30) // (bs * hkv * d * 2 * es)
    pages = [-(-kv_len // page)] * bs
    total = sum(pages)
    cache = torch.randn(total, 2, page, hkv, d, device=DEV, dtype=torch.bfloat16)
    if fp8:
        cache = cache.to(torch.float8_e4m3fn)
    indices = torch.randperm(total).to(torch.int32)
    indptr = torch.tensor([0] + list(np.cumsum(pages)), dtype=torch.int32)
    last = torch.tensor([(kv_len - 1) % page + 1] * bs, dtype=torch.int32)
    q = torch.randn(bs, hq, d, device=DEV, dtype=torch.bfloat16)
    out = torch.empty_like(q)
    w = fi.BatchDecodeWithPagedKVCacheWrapper(torch.zeros(128 << 20, dtype=torch.uint8, device=DEV), "NHD")
    w.plan(indptr, indices, last, hq, hkv, d, page, q_data_type=torch.bfloat16, kv_data_type=cache.dtype)
    for _ in range(5):
        w.run(q, cache, out=out)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    a.record()
    for _ in range(100):
        w.run(q, cache, out=out)
    b.record()
    torch.cuda.synchronize()
    us = a.elapsed_time(b) * 10
    nbytes = bs * kv_len * hkv * d * 2 * es
    ref_txt = ""
    so = Path(__file__).resolve().parent.parent / "oracle" / "_ref" / "liboracle_ref.so"
    if so.exists() and not fp8:  # the reference's own kernel at this head_dim (16-bit KV instantiations)
        import ctypes as C
        ref = C.CDLL(str(so))
        ref.ref_last_error.restype = C.c_char_p
        vp = C.c_void_p
        k, v = cache[:, 0], cache[:, 1]
        fws = torch.zeros(256 << 20, dtype=torch.uint8, device=DEV)
        iws = torch.zeros(8 << 20, dtype=torch.uint8, device=DEV)
        pws = torch.zeros(8 << 20, dtype=torch.uint8).pin_memory()
        ip_d, ix_d, lp_d = indptr.to(DEV), indices.to(DEV), last.to(DEV)
        info = torch.zeros(16, dtype=torch.int64)

        def call():
            rc = ref.ref_batch_decode(
                vp(fws.data_ptr()), C.c_size_t(fws.numel()), vp(iws.data_ptr()), vp(pws.data_ptr()),
                C.c_size_t(iws.numel()), vp(indptr.data_ptr()), bs, hq, hkv, d, page, 1, 1, 0, 0,
                vp(q.data_ptr()), C.c_int64(q.stride(0)), C.c_int64(q.stride(1)), vp(k.data_ptr()),
                vp(v.data_ptr()), C.c_int64(k.stride(0)), C.c_int64(k.stride(1)), C.c_int64(k.stride(2)),
                vp(ix_d.data_ptr()), vp(ip_d.data_ptr()), vp(lp_d.data_ptr()), vp(out.data_ptr()), None,
                C.c_float(1.0 / d ** 0.5), vp(torch.cuda.current_stream().cuda_stream), vp(info.data_ptr()))
            assert rc == 0, ref.ref_last_error()

        for _ in range(3):
            call()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()  # plan's H2D + kernels; the host-side planning stays off the timeline
        with torch.cuda.graph(g):
            call()
        a.record()
        for _ in range(20):
            g.replay()
        b.record()
        torch.cuda.synchronize()
        ref_us = a.elapsed_time(b) * 50
        ref_txt = f"   reference's own kernel: {ref_us:7.1f} us  {nbytes / ref_us / 1e6:.2f} TB/s"
    print(f"{'e4m3' if fp8 else 'bf16'} D={d} kv_len={kv_len}: {us:7.1f} us  {nbytes / us / 1e6:.2f} TB/s  "
          f"(ctas {int(w._plan_info[1])}, stages {int(w._plan_info[22])}){ref_txt}")
    del cache
