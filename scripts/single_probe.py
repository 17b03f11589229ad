"""single_decode_with_kv_cache (BASELINE configs[0] head shape: 32 / 8 heads, D 128, fp16) — ours vs the
reference's own kernel (oracle/_ref), GPU time per call over back-to-back calls."""
import ctypes as C
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import flashinfer_b200 as fi  # noqa: E402

DEV = "cuda:0"
ref = C.CDLL(str(ROOT / "oracle" / "_ref" / "liboracle_ref.so"))
ref.ref_last_error.restype = C.c_char_p
hq, hkv, d = 32, 8, 128
vp = C.c_void_p
for kv_len in (512, 2048, 8192, 32768, 131072):
    q = torch.randn(hq, d, device=DEV, dtype=torch.float16)
    k = torch.randn(kv_len, hkv, d, device=DEV, dtype=torch.float16)
    v = torch.randn(kv_len, hkv, d, device=DEV, dtype=torch.float16)
    tmp = torch.zeros(32 << 20, dtype=torch.uint8, device=DEV)
    o_ref = torch.empty_like(q)

    def ours():
        return fi.single_decode_with_kv_cache(q, k, v)

    def theirs():
        rc = ref.ref_single_decode(vp(q.data_ptr()), vp(k.data_ptr()), vp(v.data_ptr()), vp(tmp.data_ptr()),
                                   vp(o_ref.data_ptr()), None, hq, hkv, d, kv_len, 0, 0, 0,
                                   C.c_float(1.0 / d ** 0.5), vp(torch.cuda.current_stream().cuda_stream))
        assert rc == 0, ref.ref_last_error()

    res = []
    for fn in (ours, theirs):
        for _ in range(10):
            fn()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(20):
                fn()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        for _ in range(10):
            g.replay()
        b.record()
        torch.cuda.synchronize()
        res.append(a.elapsed_time(b) * 1e3 / 200)
    err = (ours().float() - o_ref.float()).abs().max().item()
    print(f"kv_len {kv_len:7d}: ours {res[0]:7.2f} us  reference {res[1]:7.2f} us  max|do| {err:.1e}")
