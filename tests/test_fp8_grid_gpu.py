"""GPU: fp8-KV decode grid over both tcgen05 variants (one KV head per unit; head-pair plans, which the
wrapper asks for on NHD caches) and the register path (pages below 8) — q dtype x KV format x layout x
page size x head shape x logits soft cap, short and split requests in one batch — against the oracle's
fp32 attention on the stored fp8 values (o <= 3e-2, lse <= 1e-3).  The reference's own
grid has e4m3 only and 4 KV heads only (tests/attention/test_batch_decode_kernels.py:58-70)."""
import zlib

import pytest
import torch

from helpers import make_paged_case, make_q
from oracle import attention_ref as orc

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
_WS = {}


def _workspace():
    if "ws" not in _WS:
        _WS["ws"] = torch.empty(128 << 20, dtype=torch.uint8, device=DEV)
    return _WS["ws"]


@pytest.mark.parametrize("q_dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("kv_dtype", [torch.float8_e4m3fn, torch.float8_e5m2])
@pytest.mark.parametrize("layout", ["NHD", "HND"])
@pytest.mark.parametrize("page", [4, 8, 16, 32])
@pytest.mark.parametrize("hq,hkv", [(8, 8), (16, 4), (32, 4), (6, 2), (8, 1), (48, 2)])
@pytest.mark.parametrize("soft_cap", [0.0, 25.0])
def test_fp8_kv_grid(q_dtype, kv_dtype, layout, page, hq, hkv, soft_cap):
    import flashinfer_b200 as fi

    d = 128
    lens = [130, 1, 6000, 0, 63, 64, 1100, 257]
    seed = zlib.crc32(repr((str(q_dtype), str(kv_dtype), layout, page, hq, hkv)).encode()) % 10007
    cache, indptr, indices, last = make_paged_case(lens, hkv, d, page, layout, torch.float32, seed=seed)
    cache = (cache * 0.5).to(kv_dtype)
    q = make_q(len(lens), hq, d, q_dtype, seed=seed + 1)
    w = fi.BatchDecodeWithPagedKVCacheWrapper(_workspace(), layout)
    w.plan(indptr, indices, last, hq, hkv, d, page, q_data_type=q_dtype, kv_data_type=kv_dtype,
           logits_soft_cap=soft_cap or None)
    pairs = layout == "NHD" and hkv % 2 == 0 and hq // hkv <= 8 and page >= 8
    assert int(w._plan_info[31]) == int(pairs)
    o, lse = w.run(q.to(DEV), cache.to(DEV), k_scale=0.5, v_scale=2.0, return_lse=True)
    o_ref, lse_ref = orc.batch_decode_ref(q, cache, indptr, indices, last, layout, k_scale=0.5, v_scale=2.0,
                                          logits_soft_cap=soft_cap)
    assert (o.float().cpu() - o_ref).abs().max().item() <= 3e-2
    assert (lse.cpu() - lse_ref).abs().max().item() <= 1e-3
