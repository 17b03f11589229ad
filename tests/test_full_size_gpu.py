"""GPU: BASELINE.json's configurations at FULL size, checked through size-independent properties
(the CPU oracle finishes a whole C2 step in ~1 s but C3 / C4 / C5 in minutes, so at full size it
is run on a sample of requests only):

  * oracle parity on sampled requests of the full batch (o 1e-2 / 3e-2 fp8, lse 1e-3);
  * placement invariance: moving every page to another physical slot (and rewriting the page
    table accordingly) leaves o / lse BIT-identical — the page-table gather has no effect on the
    arithmetic;
  * power-of-two scaling of V scales o exactly (bit-exact) and leaves lse bit-identical;
  * split consistency: decode over the first half of every request (+) decode over the second
    half, combined with merge_state, equals the full decode (lse 1e-3) — the attention-state
    algebra of the reference (state.cuh:53-62) at full size;
  * plan independence: disable_split_kv (no unit is ever cut) agrees with the default plan.

Configs: C2 (bs64 x kv4096, bf16), C3 (bs128 x kv8192, e4m3 KV; all 8 KV heads on one GPU and the
Hkv = 1 per-rank shape), C4 (bs256, log-uniform 512..32768), C5 per-rank shape (bs16 x kv131072).
"""
import math

import numpy as np
import pytest
import torch

from oracle import attention_ref as orc

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
PAGE, D = 16, 128


def _fi():
    import flashinfer_b200 as fi

    return fi


def _lens(name):
    if name == "c2":
        return [4096] * 64, 32, 8, torch.bfloat16
    if name == "c3":
        return [8192] * 128, 64, 8, torch.float8_e4m3fn
    if name == "c3r":
        return [8192] * 128, 8, 1, torch.float8_e4m3fn
    if name == "c4":
        rng = np.random.default_rng(0)
        lens = np.exp(rng.uniform(math.log(512), math.log(32768), 256)).round().astype(int).tolist()
        return lens, 32, 8, torch.bfloat16
    if name == "c5r":
        return [131072] * 16, 8, 1, torch.bfloat16
    raise KeyError(name)


def _table(kv_lens, seed):
    pages = [-(-l // PAGE) for l in kv_lens]
    total = sum(pages)
    g = torch.Generator().manual_seed(seed)
    indices = torch.randperm(total, generator=g).to(torch.int32)
    indptr = torch.zeros(len(kv_lens) + 1, dtype=torch.int32)
    indptr[1:] = torch.tensor(np.cumsum(pages), dtype=torch.int32)
    last = torch.tensor([((l - 1) % PAGE + 1) for l in kv_lens], dtype=torch.int32)
    return indptr, indices, last, total


def _case(name):
    kv_lens, hq, hkv, kv_dtype = _lens(name)
    indptr, indices, last, total = _table(kv_lens, seed=3)
    g = torch.Generator(device=DEV).manual_seed(11)
    cache = torch.randn(total, 2, PAGE, hkv, D, device=DEV, dtype=torch.bfloat16, generator=g)
    scales = {}
    if kv_dtype == torch.float8_e4m3fn:
        # per-tensor scales as the reference's calibration test (test_decode_fp8_calibration_scale.py:131-137)
        k_scale = float(cache[:, 0].float().abs().amax() / 256)
        v_scale = float(cache[:, 1].float().abs().amax() / 256)
        cache[:, 0] /= k_scale
        cache[:, 1] /= v_scale
        cache = cache.to(kv_dtype)
        scales = dict(k_scale=k_scale, v_scale=v_scale)
    q = torch.randn(len(kv_lens), hq, D, device=DEV, dtype=torch.bfloat16, generator=g)
    return kv_lens, hq, hkv, kv_dtype, indptr, indices, last, cache, q, scales


def _bytes(t):
    """same storage viewed as an integer type (index ops on float8 tensors are not implemented)"""
    return t.view(torch.uint8) if t.element_size() == 1 else t.view(torch.int16)


def _decode(cache, q, indptr, indices, last, hq, hkv, kv_dtype, scales, **plan_kw):
    fi = _fi()
    w = fi.BatchDecodeWithPagedKVCacheWrapper(torch.zeros(256 << 20, dtype=torch.uint8, device=DEV), "NHD")
    w.plan(indptr, indices, last, hq, hkv, D, PAGE, q_data_type=torch.bfloat16, kv_data_type=kv_dtype,
           **plan_kw)
    o, lse = w.run(q, cache, return_lse=True, **scales)
    torch.cuda.synchronize()
    return o, lse, w


@pytest.mark.parametrize("name", ["c2", "c3", "c3r", "c4", "c5r"])
def test_full_size_sampled_oracle_and_placement_invariance(name):
    kv_lens, hq, hkv, kv_dtype, indptr, indices, last, cache, q, scales = _case(name)
    o, lse, w = _decode(cache, q, indptr, indices, last, hq, hkv, kv_dtype, scales)
    assert torch.isfinite(o.float()).all() and torch.isfinite(lse).all()

    # (1) oracle on sampled requests: shortest, longest, first, last
    bs = len(kv_lens)
    sample = sorted({0, bs - 1, int(np.argmin(kv_lens)), int(np.argmax(kv_lens))})
    if name == "c5r":
        sample = sample[:1]  # 131072 tokens each: one is enough for the CPU
    k4, v4 = cache[:, 0], cache[:, 1]
    for i in sample:
        b, e = int(indptr[i]), int(indptr[i + 1])
        pages = indices[b:e].long().to(DEV)
        k = _bytes(k4)[pages].view(cache.dtype).reshape(-1, hkv, D)[: kv_lens[i]].float().cpu()
        v = _bytes(v4)[pages].view(cache.dtype).reshape(-1, hkv, D)[: kv_lens[i]].float().cpu()
        o_ref, lse_ref = orc.decode_attention_ref(q[i].cpu(), k, v,
                                                  sm_scale=scales.get("k_scale", 1.0) / math.sqrt(D))
        o_ref = o_ref * scales.get("v_scale", 1.0)
        tol = 3e-2 if kv_dtype == torch.float8_e4m3fn else 1e-2
        assert (o[i].float().cpu() - o_ref).abs().max().item() <= tol, (name, i)
        assert (lse[i].cpu() - lse_ref).abs().max().item() <= 1e-3, (name, i)

    # (2) placement invariance: page j moves to slot perm[j]; same logical KV, same plan -> same bits
    total = cache.shape[0]
    g = torch.Generator().manual_seed(99)
    perm = torch.randperm(total, generator=g)
    moved = torch.empty_like(cache)
    _bytes(moved)[perm.to(DEV)] = _bytes(cache)
    indices2 = perm[indices.long()].to(torch.int32)
    o2, lse2, _ = _decode(moved, q, indptr, indices2, last, hq, hkv, kv_dtype, scales)
    assert torch.equal(o2, o) and torch.equal(lse2, lse)
    del moved


@pytest.mark.parametrize("name", ["c2", "c3", "c4"])
def test_full_size_v_scaling_is_exact(name):
    kv_lens, hq, hkv, kv_dtype, indptr, indices, last, cache, q, scales = _case(name)
    o, lse, _ = _decode(cache, q, indptr, indices, last, hq, hkv, kv_dtype, scales)
    if kv_dtype == torch.float8_e4m3fn:
        # stored bytes stay, the per-tensor scale doubles (fused into the epilogue)
        s2 = dict(scales, v_scale=scales["v_scale"] * 2.0)
        o2, lse2, _ = _decode(cache, q, indptr, indices, last, hq, hkv, kv_dtype, s2)
    else:
        c2 = cache.clone()
        c2[:, 1] *= 2  # exact in bf16
        o2, lse2, _ = _decode(c2, q, indptr, indices, last, hq, hkv, kv_dtype, scales)
        del c2
    assert torch.equal(lse2, lse)
    assert torch.equal(o2.float(), o.float() * 2)


@pytest.mark.parametrize("name", ["c2", "c3", "c4"])
def test_full_size_split_consistency_and_plan_independence(name):
    fi = _fi()
    kv_lens, hq, hkv, kv_dtype, indptr, indices, last, cache, q, scales = _case(name)
    o, lse, w = _decode(cache, q, indptr, indices, last, hq, hkv, kv_dtype, scales)

    # first half / second half of every request on page boundaries, merged with merge_state
    npages = (indptr[1:] - indptr[:-1]).tolist()
    h1 = [n // 2 for n in npages]
    ip1 = torch.zeros_like(indptr)
    ip2 = torch.zeros_like(indptr)
    ip1[1:] = torch.tensor(np.cumsum(h1), dtype=torch.int32)
    ip2[1:] = torch.tensor(np.cumsum([n - a for n, a in zip(npages, h1)]), dtype=torch.int32)
    ix1 = torch.cat([indices[int(indptr[i]): int(indptr[i]) + h1[i]] for i in range(len(kv_lens))])
    ix2 = torch.cat([indices[int(indptr[i]) + h1[i]: int(indptr[i + 1])] for i in range(len(kv_lens))])
    full = torch.full_like(last, PAGE)
    o1, l1, _ = _decode(cache, q, ip1, ix1, full, hq, hkv, kv_dtype, scales)
    o2, l2, _ = _decode(cache, q, ip2, ix2, last, hq, hkv, kv_dtype, scales)
    om, lm = fi.merge_state(o1, l1, o2, l2)
    tol = 3e-2 if kv_dtype == torch.float8_e4m3fn else 1e-2
    assert (om.float() - o.float()).abs().max().item() <= tol
    assert (lm - lse).abs().max().item() <= 1e-3

    # a plan that never cuts a unit (different work list, no partials, no merge)
    o3, l3, w3 = _decode(cache, q, indptr, indices, last, hq, hkv, kv_dtype, scales, disable_split_kv=True)
    assert int(w3._plan_info[3]) == 0
    if name == "c4":  # the uniform C2 / C3 plans hold whole requests per head group: nothing to merge either
        assert int(w._plan_info[3]) > 0
    assert (o3.float() - o.float()).abs().max().item() <= tol
    assert (l3 - lse).abs().max().item() <= 1e-3
