"""Comparator: time the REFERENCE's own CUDA-core decode kernel (oracle/_ref/liboracle_ref.so,
compiled from /root/reference headers) on the BASELINE configs, same box, same inputs shape.
python scripts/ref_compare.py [c2|c3|c3r|c4|c5r ...]   -> one JSON line per config."""
import ctypes as C
import json
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "scripts"))
from perf_probe import config  # noqa: E402

DEV = "cuda:0"
DT = {torch.float16: 0, torch.bfloat16: 1, torch.float8_e4m3fn: 2}


def main():
    ref = C.CDLL(str(ROOT / "oracle" / "_ref" / "liboracle_ref.so"))
    ref.ref_last_error.restype = C.c_char_p
    names = sys.argv[1:] or ["c2"]
    for name in names:
        c = config(name)
        page, d = 16, 128
        kv_lens, hq, hkv, kv_dtype = c["kv_lens"], c["hq"], c["hkv"], c["kv_dtype"]
        pages = [-(-l // page) for l in kv_lens]
        total = sum(pages)
        es = 1 if kv_dtype == torch.float8_e4m3fn else 2
        cache = torch.randn(total, 2, page, hkv, d, device=DEV, dtype=torch.bfloat16)
        cache = cache.to(kv_dtype) if es == 1 else cache
        indices = torch.randperm(total, device=DEV).to(torch.int32)
        indptr = torch.tensor([0] + list(np.cumsum(pages)), dtype=torch.int32)
        last = torch.tensor([((l - 1) % page + 1) for l in kv_lens], dtype=torch.int32)
        q = torch.randn(len(kv_lens), hq, d, device=DEV, dtype=torch.bfloat16)
        k, v = cache[:, 0], cache[:, 1]
        fws = torch.zeros(256 << 20, dtype=torch.uint8, device=DEV)
        iws = torch.zeros(8 << 20, dtype=torch.uint8, device=DEV)
        pws = torch.zeros(8 << 20, dtype=torch.uint8).pin_memory()
        o = torch.empty_like(q)
        ip_d, lp_d = indptr.to(DEV), last.to(DEV)
        info = torch.zeros(16, dtype=torch.int64)
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)
        vp = C.c_void_p

        def call():
            rc = ref.ref_batch_decode(
                vp(fws.data_ptr()), C.c_size_t(fws.numel()), vp(iws.data_ptr()), vp(pws.data_ptr()),
                C.c_size_t(iws.numel()), vp(indptr.data_ptr()), len(kv_lens), hq, hkv, d, page, 1,
                DT[kv_dtype], 0, 0, vp(q.data_ptr()), C.c_int64(q.stride(0)), C.c_int64(q.stride(1)),
                vp(k.data_ptr()), vp(v.data_ptr()), C.c_int64(k.stride(0)), C.c_int64(k.stride(1)),
                C.c_int64(k.stride(2)), vp(indices.data_ptr()), vp(ip_d.data_ptr()), vp(lp_d.data_ptr()),
                vp(o.data_ptr()), None, C.c_float(1.0 / d ** 0.5),
                vp(torch.cuda.current_stream().cuda_stream), vp(info.data_ptr()))
            assert rc == 0, ref.ref_last_error()

        for _ in range(3):
            call()
        torch.cuda.synchronize()
        # capture plan's H2D copy + kernel(s) in a CUDA graph so that the host-side planning
        # time is not on the measured timeline (the reference benchmarks with graphs too)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            call()
        ts = []
        for _ in range(15):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            graph.replay()
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3)
        nbytes = sum(kv_lens) * hkv * d * 2 * es + len(kv_lens) * hq * d * 4
        us = float(np.median(ts))
        print(json.dumps({"config": name, "impl": "reference CUDA-core kernel (decode.cuh, sm_100a build)",
                          "us_per_step": us, "tbps": nbytes / us / 1e6,
                          "split_kv": bool(info[9]), "padded_batch_size": int(info[0])}))


if __name__ == "__main__":
    main()
