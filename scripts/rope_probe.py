"""ROPE_LLAMA decode at the C2 shape: ours (in-register K rotation in the MMA kernel; set
FI_B200_ROPE_CUDA_CORE=1 for the older CUDA-core variant kernel) and the reference's own kernel
instantiated with PosEncodingMode::kRoPELlama (oracle/_ref).  python scripts/rope_probe.py [ours|ref]"""
import ctypes as C
import json
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import flashinfer_b200 as fi  # noqa: E402
from bench import HEAD_DIM, PAGE, page_table, workload  # noqa: E402

DEV = torch.device("cuda", 0)


def timed(fn, n=30):
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)
    for _ in range(3):
        fn()
    ts = []
    for _ in range(n):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        b.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    return float(np.median(ts))


def main():
    which = sys.argv[1] if len(sys.argv) > 1 else "ours"
    wl = workload(sys.argv[2] if len(sys.argv) > 2 else "c2")
    kv_lens, hq, hkv = wl["kv_lens"], wl["hq"], wl["hkv"]
    bs = len(kv_lens)
    indptr, indices, last, total = page_table(kv_lens)
    cache = torch.randn(total, 2, PAGE, hkv, HEAD_DIM, device=DEV, dtype=torch.bfloat16)
    q = torch.randn(bs, hq, HEAD_DIM, device=DEV, dtype=torch.bfloat16)
    out = torch.empty_like(q)
    nbytes = sum(kv_lens) * hkv * HEAD_DIM * 4 + bs * hq * HEAD_DIM * 4
    if which == "ours":
        ws = torch.zeros(128 << 20, dtype=torch.uint8, device=DEV)
        for mode in ("NONE", "ROPE_LLAMA"):
            w = fi.BatchDecodeWithPagedKVCacheWrapper(ws, "NHD")
            w.plan(indptr, indices, last, hq, hkv, HEAD_DIM, PAGE, pos_encoding_mode=mode,
                   q_data_type=torch.bfloat16, rope_theta=5e5)
            us = timed(lambda: w.run(q, cache, out=out))
            print(json.dumps({"impl": "ours", "mode": mode, "us": us, "tbps": nbytes / us / 1e6}))
    else:
        ref = C.CDLL(str(ROOT / "oracle" / "_ref" / "liboracle_ref.so"))
        ref.ref_last_error.restype = C.c_char_p
        k, v = cache[:, 0], cache[:, 1]
        fws = torch.zeros(256 << 20, dtype=torch.uint8, device=DEV)
        iws = torch.zeros(8 << 20, dtype=torch.uint8, device=DEV)
        pws = torch.zeros(8 << 20, dtype=torch.uint8).pin_memory()
        ip_d, ix_d, lp_d = indptr.to(DEV), indices.to(DEV), last.to(DEV)
        info = torch.zeros(16, dtype=torch.int64)
        vp = C.c_void_p

        def call(mode):
            rc = ref.ref_batch_decode_rope(
                vp(fws.data_ptr()), C.c_size_t(fws.numel()), vp(iws.data_ptr()), vp(pws.data_ptr()),
                C.c_size_t(iws.numel()), vp(indptr.data_ptr()), bs, hq, hkv, HEAD_DIM, PAGE, 1, 1, 0, 0,
                vp(q.data_ptr()), C.c_int64(q.stride(0)), C.c_int64(q.stride(1)),
                vp(k.data_ptr()), vp(v.data_ptr()), C.c_int64(k.stride(0)), C.c_int64(k.stride(1)),
                C.c_int64(k.stride(2)), vp(ix_d.data_ptr()), vp(ip_d.data_ptr()), vp(lp_d.data_ptr()),
                vp(out.data_ptr()), None, C.c_float(1.0 / HEAD_DIM ** 0.5), C.c_float(1.0),
                C.c_float(5e5), vp(torch.cuda.current_stream().cuda_stream), vp(info.data_ptr()), mode)
            assert rc == 0, ref.ref_last_error()

        call(1)  # plan once (as the reference's wrapper), then time run only
        us = timed(lambda: call(2))
        print(json.dumps({"impl": "reference kernel, PosEncodingMode::kRoPELlama", "mode": "ROPE_LLAMA",
                          "us": us, "tbps": nbytes / us / 1e6}))


if __name__ == "__main__":
    main()
