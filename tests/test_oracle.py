"""CPU: the oracle restatement (oracle/attention_ref.py) against the golden vectors produced
by the reference's own test oracle, plus internal consistency of the merge restatements."""
import math
from pathlib import Path

import numpy as np
import pytest
import torch

from helpers import golden_cases, golden_inputs, make_paged_case, make_q
from oracle import attention_ref as orc

Z, NAMES = golden_cases()


@pytest.mark.parametrize("name", NAMES)
def test_oracle_matches_reference_golden(name):
    q, k, v, ok = golden_inputs(Z, name)
    if not ok:
        pytest.skip("RNG of this machine does not reproduce the un-stored golden inputs")
    o, lse2 = orc.decode_attention_ref(q, k, v)
    # the reference oracle rounds o to q's dtype; o_fp32 is the same function on fp32 inputs
    np.testing.assert_allclose(o.numpy(), Z[f"{name}/o_fp32"], atol=2e-6, rtol=1e-5)
    np.testing.assert_allclose(o.to(q.dtype).float().numpy(), Z[f"{name}/o"], atol=4e-3, rtol=1e-2)
    np.testing.assert_allclose(lse2.numpy(), Z[f"{name}/lse2"], atol=1e-5, rtol=1e-6)


def test_paged_gather_equals_contiguous():
    kv_lens = [1, 54, 97, 16, 0, 33]
    for layout in ("NHD", "HND"):
        cache, indptr, indices, last = make_paged_case(kv_lens, 2, 64, 8, layout, torch.float16)
        kc, vc = orc.unpack_cache(cache, layout)
        for i, L in enumerate(kv_lens):
            k, v = orc.gather_paged_kv(kc, vc, indptr, indices, last, i, layout)
            assert k.shape == (L, 2, 64) and v.shape == (L, 2, 64)
            for t in (0, L // 2, L - 1):
                if L == 0:
                    continue
                page = int(indices[int(indptr[i]) + t // 8])
                row = t % 8
                exp = cache[page, 0, row] if layout == "NHD" else cache[page, 0, :, row]
                assert torch.equal(k[t], exp)


def test_empty_request_sentinel():
    cache, indptr, indices, last = make_paged_case([0, 9], 1, 64, 16, "HND", torch.float16)
    q = make_q(2, 4, 64, torch.float16)
    o, lse = orc.batch_decode_ref(q, cache, indptr, indices, last, "HND")
    assert torch.all(o[0] == 0) and torch.all(lse[0] == orc.NEG_INF_SENTINEL)
    assert torch.isfinite(lse[1]).all()


def test_split_then_merge_equals_whole():
    """merge_states over chunked attention == attention over the whole KV (the identity the
    split-KV kernel relies on, reference docs/tutorials/recursive_attention.rst:18-36)."""
    g = torch.Generator().manual_seed(3)
    q = torch.randn(8, 64, generator=g)
    k = torch.randn(300, 2, 64, generator=g)
    v = torch.randn(300, 2, 64, generator=g)
    o, lse = orc.decode_attention_ref(q, k, v)
    cuts = [0, 64, 128, 299, 300]
    parts = [orc.decode_attention_ref(q, k[a:b], v[a:b]) for a, b in zip(cuts[:-1], cuts[1:])]
    vs = torch.stack([p[0] for p in parts]).unsqueeze(0)  # [1, n, H, D]
    ss = torch.stack([p[1] for p in parts]).unsqueeze(0)
    vm, sm = orc.merge_states_ref(vs, ss)
    torch.testing.assert_close(vm[0], o, atol=1e-5, rtol=1e-5)
    torch.testing.assert_close(sm[0], lse, atol=1e-5, rtol=1e-5)
    # 2-way and variable-length forms agree with the n-way one
    v2, s2 = orc.merge_state_ref(vs[:, 0], ss[:, 0], vs[:, 1], ss[:, 1])
    v2n, s2n = orc.merge_states_ref(vs[:, :2], ss[:, :2])
    torch.testing.assert_close(v2, v2n, atol=1e-6, rtol=1e-6)
    torch.testing.assert_close(s2, s2n, atol=1e-6, rtol=1e-6)
    vv, sv = orc.variable_length_merge_states_ref(vs[0], ss[0], [0, 4, 4])
    torch.testing.assert_close(vv[0], o, atol=1e-5, rtol=1e-5)
    assert torch.all(vv[1] == 0) and torch.all(sv[1] == orc.NEG_INF_SENTINEL)


def test_soft_cap_and_window():
    g = torch.Generator().manual_seed(4)
    q = torch.randn(4, 64, generator=g)
    k = torch.randn(50, 4, 64, generator=g)
    v = torch.randn(50, 4, 64, generator=g)
    o_w, _ = orc.decode_attention_ref(q, k, v, window_left=9)
    o_s, _ = orc.decode_attention_ref(q, k[40:], v[40:])
    torch.testing.assert_close(o_w, o_s)
    o_c, _ = orc.decode_attention_ref(q, k, v, logits_soft_cap=30.0)
    logits = 30.0 * torch.tanh(torch.einsum("hd,nhd->hn", q, k) / math.sqrt(64) / 30.0)
    torch.testing.assert_close(o_c, torch.einsum("hn,nhd->hd", torch.softmax(logits, -1), v))


def test_reference_uniform_chunking_covers_everything():
    pages = [256] * 64
    split, chunk, req, tile = orc.reference_uniform_chunking(pages, 8, 2368, 16)
    assert split and chunk == 1024 and len(req) == 256  # SURVEY.md §8(a3) C2 figure
    split, chunk, req, tile = orc.reference_uniform_chunking([4] * 512, 8, 2368, 16)
    assert not split and len(req) == 512


# ---------------------------------------------------------------- ALiBi
def _alibi_cases():
    z = np.load(Path(__file__).resolve().parent / "golden" / "alibi_golden.npz")
    return z, [str(n) for n in z["names"]]


def _alibi_inputs(z, name):
    H, D, L, is_bf16 = [int(x) for x in z[f"{name}/meta"]]
    dt = torch.bfloat16 if is_bf16 else torch.float16
    q, k, v, o = (torch.from_numpy(z[f"{name}/{key}"].copy()).view(dt) for key in ("q", "k", "v", "o"))
    return q, k, v, o, torch.from_numpy(z[f"{name}/slopes"].copy())


@pytest.mark.parametrize("name", _alibi_cases()[1])
def test_oracle_alibi_matches_reference_alibi_attention(name):
    """oracle ALiBi == the reference's own alibi_attention (tests/test_helpers/alibi_reference.py),
    executed in the build container by tests/golden/make_golden_alibi.py; slopes == its get_slopes."""
    from flashinfer_b200.utils import get_alibi_slopes

    z, _ = _alibi_cases()
    q, k, v, o_gold, slopes = _alibi_inputs(z, name)
    assert torch.equal(get_alibi_slopes(q.shape[0]), slopes)
    o, _ = orc.decode_attention_ref(q, k, v, alibi_slopes=slopes)
    # the golden output was rounded to q's dtype by the reference function
    assert torch.equal(o.to(q.dtype), o_gold) or (o - o_gold.float()).abs().max().item() <= 2e-3


# ---------------------------------------------------------------- RoPE on the fly
def _rope_cases():
    z = np.load(Path(__file__).resolve().parent / "golden" / "rope_golden.npz")
    return z, [str(n) for n in z["names"]]


def rope_inputs(z, name):
    Hq, Hkv, D, L, is_bf16 = [int(x) for x in z[f"{name}/meta"]]
    dt = torch.bfloat16 if is_bf16 else torch.float16
    q, k, v = (torch.from_numpy(z[f"{name}/{key}"].copy()).view(dt) for key in ("q", "k", "v"))
    return q, k, v, float(z[f"{name}/theta"][0]), torch.from_numpy(z[f"{name}/o"].copy()), \
        torch.from_numpy(z[f"{name}/lse2"].copy())


@pytest.mark.parametrize("name", _rope_cases()[1])
def test_oracle_rope_matches_reference_helpers(name):
    """oracle ROPE_LLAMA == the reference's rope_reference.py rotation + its attention_ref, executed
    in the build container by tests/golden/make_golden_rope.py"""
    z, _ = _rope_cases()
    q, k, v, theta, o_gold, lse_gold = rope_inputs(z, name)
    o, lse = orc.decode_attention_ref(q, k, v, rope=(1.0, theta))
    assert (o - o_gold).abs().max().item() <= 2e-5
    assert (lse - lse_gold).abs().max().item() <= 2e-4


# ---------------------------------------------------------------- attention sinks
def _sink_cases():
    z = np.load(Path(__file__).resolve().parent / "golden" / "sink_golden.npz")
    return z, [str(n) for n in z["names"]]


@pytest.mark.parametrize("name", _sink_cases()[1])
def test_oracle_sinks_match_reference_sink_attention(name):
    """oracle sinks == the reference's own sink_attention_unified (incremental mode, built on its
    sink_softmax, tests/test_helpers/sink_attention_reference.py), executed in the build container
    by tests/golden/make_golden_sink.py."""
    z, _ = _sink_cases()
    B, L, Hq, Hkv, D, win, is_bf16 = [int(x) for x in z[f"{name}/meta"]]
    dt = torch.bfloat16 if is_bf16 else torch.float16
    q, k, v, o_gold = (torch.from_numpy(z[f"{name}/{key}"].copy()).view(dt) for key in ("q", "k", "v", "o"))
    q, o_gold = q.view(B, Hq, D), o_gold.view(B, Hq, D)
    k, v = k.view(B, L, Hkv, D), v.view(B, L, Hkv, D)
    sink = torch.from_numpy(z[f"{name}/sink"].copy())
    for b in range(B):
        o, _ = orc.decode_attention_ref(q[b], k[b], v[b], window_left=win, sinks=sink)
        assert (o - o_gold[b].float()).abs().max().item() <= 2e-3  # the golden output is rounded to q's dtype
