"""Page-table maintenance ops (flashinfer_b200/page.py, csrc/page.cu) against the CPU oracle
(oracle/page_ref.py) and the known answers of the reference's docstrings.  Byte work: bit-exact."""
import numpy as np
import pytest
import torch

from oracle import page_ref

DEV = "cuda:0"

# flashinfer/page.py:190-199 (docstring of get_batch_indices_positions)
DOC1 = dict(indptr=[0, 1, 3, 6, 10], seq_lens=[5, 5, 5, 5], nnz=10,
            batch_indices=[0, 1, 1, 2, 2, 2, 3, 3, 3, 3], positions=[4, 3, 4, 2, 3, 4, 1, 2, 3, 4])
# flashinfer/page.py:352-383 (docstring of append_paged_kv_cache): append lengths 45 + 8 + 25 + 22
DOC2_LENS = [45, 8, 25, 22]


def _doc2():
    indptr = np.concatenate([[0], np.cumsum(DOC2_LENS)])
    pages = np.array([3, 1, 2, 2])
    last = np.array([13, 8, 9, 6])
    seq_lens = (pages - 1) * 16 + last
    bi = np.concatenate([np.full(n, i) for i, n in enumerate(DOC2_LENS)])
    pos = np.concatenate([np.arange(n) for n in DOC2_LENS])
    return indptr, seq_lens, bi, pos, pages, last


def test_oracle_matches_reference_docstrings():
    b, p = page_ref.get_batch_indices_positions_ref(DOC1["indptr"], DOC1["seq_lens"], DOC1["nnz"])
    assert b.tolist() == DOC1["batch_indices"] and p.tolist() == DOC1["positions"]
    indptr, seq_lens, bi, pos, _, _ = _doc2()
    b, p = page_ref.get_batch_indices_positions_ref(indptr, seq_lens, 100)
    assert np.array_equal(b, bi) and np.array_equal(p, pos)


def test_oracle_append_places_rows():
    g = torch.Generator().manual_seed(0)
    cache = torch.zeros(6, 2, 4, 2, 8)
    k, v = torch.randn(5, 2, 8, generator=g), torch.randn(5, 2, 8, generator=g)
    # request 0: pages [4, 1], tokens 2..5 appended (positions 2,3,4,5); request 1: page [3], position 0
    page_ref.append_paged_kv_cache_ref(k, v, [0, 0, 0, 0, 1], [2, 3, 4, 5, 0], cache,
                                       [4, 1, 3], [0, 2, 3], [2, 1], "NHD")
    assert torch.equal(cache[4, 0, 2], k[0]) and torch.equal(cache[4, 1, 3], v[1])
    assert torch.equal(cache[1, 0, 0], k[2]) and torch.equal(cache[1, 1, 1], v[3])
    assert torch.equal(cache[3, 0, 0], k[4])
    assert cache[0].abs().sum() == 0 and cache[4, :, :2].abs().sum() == 0


@pytest.mark.gpu
def test_get_batch_indices_positions_known_answers():
    import flashinfer_b200 as fi

    ip = torch.tensor(DOC1["indptr"], dtype=torch.int32, device=DEV)
    b, p = fi.get_batch_indices_positions(ip, torch.tensor(DOC1["seq_lens"]), DOC1["nnz"])  # CPU seq_lens as in the doc
    assert b.dtype == torch.int32 and b.tolist() == DOC1["batch_indices"] and p.tolist() == DOC1["positions"]
    indptr, seq_lens, bi, pos, pages, last = _doc2()
    kv_indptr = torch.tensor(np.concatenate([[0], np.cumsum(pages)]), dtype=torch.int32, device=DEV)
    sl = fi.get_seq_lens(kv_indptr, torch.tensor(last, dtype=torch.int32, device=DEV), 16)
    b, p = fi.get_batch_indices_positions(torch.tensor(indptr, dtype=torch.int32, device=DEV), sl, 100)
    assert np.array_equal(b.cpu().numpy(), bi) and np.array_equal(p.cpu().numpy(), pos)


@pytest.mark.gpu
@pytest.mark.parametrize("bs,max_new", [(1, 1), (7, 1), (300, 5), (64, 700), (5, 0)])
def test_get_batch_indices_positions_random(bs, max_new):
    import flashinfer_b200 as fi

    rng = np.random.default_rng(bs * 31 + max_new)
    new = rng.integers(0, max_new + 1, bs)
    indptr = np.concatenate([[0], np.cumsum(new)]).astype(np.int32)
    seq_lens = (new + rng.integers(0, 5000, bs)).astype(np.int32)
    nnz = int(indptr[-1])
    b, p = fi.get_batch_indices_positions(torch.tensor(indptr, device=DEV), torch.tensor(seq_lens, device=DEV), nnz)
    rb, rp = page_ref.get_batch_indices_positions_ref(indptr, seq_lens, nnz)
    assert np.array_equal(b.cpu().numpy(), rb) and np.array_equal(p.cpu().numpy(), rp)


def _append_case(layout, dtype, tuple_cache, contiguous, hkv, d, page, seed, new_lens, old_lens):
    """Build a cache that already holds old_lens tokens per request, append new_lens more."""
    g = torch.Generator().manual_seed(seed)
    bs = len(new_lens)
    tot = [o + n for o, n in zip(old_lens, new_lens)]
    pages = [-(-t // page) for t in tot]
    num_pages = sum(pages) + 5
    shape = (num_pages, 2, page, hkv, d) if layout == "NHD" else (num_pages, 2, hkv, page, d)
    cache = torch.randn(*shape, generator=g).to(dtype)
    perm = torch.randperm(num_pages, generator=g)[: sum(pages)].to(torch.int32)
    kv_indptr = torch.tensor(np.concatenate([[0], np.cumsum(pages)]), dtype=torch.int32)
    last = torch.tensor([((t - 1) % page + 1) if t > 0 else 0 for t in tot], dtype=torch.int32)
    nnz = sum(new_lens)
    if contiguous:
        k = torch.randn(nnz, hkv, d, generator=g).to(dtype)
        v = torch.randn(nnz, hkv, d, generator=g).to(dtype)
    else:  # the reference test's interleaved [nnz, 2, H, D] source (tests/attention/test_page.py:17-19)
        kv = torch.randn(nnz, 2, hkv, d, generator=g).to(dtype)
        k, v = kv[:, 0], kv[:, 1]
    append_indptr = torch.tensor(np.concatenate([[0], np.cumsum(new_lens)]), dtype=torch.int32)
    return cache, perm, kv_indptr, last, k, v, append_indptr, torch.tensor(tot, dtype=torch.int32), nnz


@pytest.mark.gpu
@pytest.mark.parametrize("layout", ["NHD", "HND"])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16, torch.float8_e4m3fn])
@pytest.mark.parametrize("tuple_cache,contiguous", [(False, True), (True, False), (False, False)])
def test_append_paged_kv_cache_bit_exact(layout, dtype, tuple_cache, contiguous):
    import flashinfer_b200 as fi

    new_lens = [45, 8, 25, 22, 0, 1]          # the reference docstring's lengths + edge cases
    old_lens = [0, 100, 7, 16, 33, 15]
    cache, indices, kv_indptr, last, k, v, a_indptr, seq_lens, nnz = _append_case(
        layout, dtype, tuple_cache, contiguous, 4, 128, 16, 11, new_lens, old_lens)
    ref = cache.clone()
    rb, rp = page_ref.get_batch_indices_positions_ref(a_indptr, seq_lens, nnz)
    page_ref.append_paged_kv_cache_ref(k, v, rb, rp, ref, indices, kv_indptr, last, layout)
    cache_d = cache.to(DEV)
    b, p = fi.get_batch_indices_positions(a_indptr.to(DEV), seq_lens.to(DEV), nnz)
    pkv = (cache_d[:, 0], cache_d[:, 1]) if tuple_cache else cache_d
    out = fi.append_paged_kv_cache(k.to(DEV) if contiguous else torch.stack([k, v], 1).to(DEV)[:, 0],
                                   v.to(DEV) if contiguous else torch.stack([k, v], 1).to(DEV)[:, 1],
                                   b, p, pkv, indices.to(DEV), kv_indptr.to(DEV), last.to(DEV), layout)
    assert out is None
    torch.cuda.synchronize()
    got = cache_d.cpu()
    assert torch.equal(got.view(torch.uint8), ref.view(torch.uint8))
    assert not torch.equal(got.view(torch.uint8), cache.view(torch.uint8))  # something was written


@pytest.mark.gpu
@pytest.mark.parametrize("d,page,hkv", [(64, 1, 1), (256, 5, 2), (80, 8, 3), (128, 32, 8)])
def test_append_shapes_and_unaligned_rows(d, page, hkv):
    """Row sizes / strides that force the narrower copy vectors (80 x 1 byte rows, odd offsets)."""
    import flashinfer_b200 as fi

    for dtype in (torch.float16, torch.float8_e5m2):
        new_lens, old_lens = [3, 0, 17, 1], [2, 9, 0, 40]
        cache, indices, kv_indptr, last, k, v, a_indptr, seq_lens, nnz = _append_case(
            "NHD", dtype, False, True, hkv, d, page, d + page, new_lens, old_lens)
        ref = cache.clone()
        rb, rp = page_ref.get_batch_indices_positions_ref(a_indptr, seq_lens, nnz)
        page_ref.append_paged_kv_cache_ref(k, v, rb, rp, ref, indices, kv_indptr, last, "NHD")
        cache_d = cache.to(DEV)
        # an element-offset view of a larger buffer: base pointers no longer 16-byte aligned
        kbuf = torch.zeros(nnz * hkv * d + 1, dtype=dtype, device=DEV)
        kd = kbuf[1:].view(nnz, hkv, d)
        kd.copy_(k)
        fi.append_paged_kv_cache(kd, v.to(DEV), torch.tensor(rb, device=DEV), torch.tensor(rp, device=DEV),
                                 cache_d, indices.to(DEV), kv_indptr.to(DEV), last.to(DEV))
        torch.cuda.synchronize()
        assert torch.equal(cache_d.cpu().view(torch.uint8), ref.view(torch.uint8))


@pytest.mark.gpu
def test_append_then_decode_serving_step():
    """One serving iteration end to end: append the new token of every request, then decode over
    the updated cache; the result must match the oracle on the oracle-updated cache."""
    import flashinfer_b200 as fi
    from oracle import attention_ref as orc

    hq, hkv, d, page = 32, 8, 128, 16
    old_lens = [54, 97, 512, 15, 16, 2047]
    new_lens = [1] * len(old_lens)
    cache, indices, kv_indptr, last, k, v, a_indptr, seq_lens, nnz = _append_case(
        "NHD", torch.bfloat16, False, True, hkv, d, page, 5, new_lens, old_lens)
    ref = cache.clone()
    rb, rp = page_ref.get_batch_indices_positions_ref(a_indptr, seq_lens, nnz)
    page_ref.append_paged_kv_cache_ref(k, v, rb, rp, ref, indices, kv_indptr, last, "NHD")
    g = torch.Generator().manual_seed(9)
    q = torch.randn(len(old_lens), hq, d, generator=g).to(torch.bfloat16)
    cache_d = cache.to(DEV)
    b, p = fi.get_batch_indices_positions(a_indptr.to(DEV), seq_lens.to(DEV), nnz)
    fi.append_paged_kv_cache(k.to(DEV), v.to(DEV), b, p, cache_d, indices.to(DEV), kv_indptr.to(DEV), last.to(DEV))
    w = fi.BatchDecodeWithPagedKVCacheWrapper(torch.zeros(64 << 20, dtype=torch.uint8, device=DEV), "NHD")
    w.plan(kv_indptr, indices, last, hq, hkv, d, page, q_data_type=torch.bfloat16)
    o, lse = w.run(q.to(DEV), cache_d, return_lse=True)
    o_ref, lse_ref = orc.batch_decode_ref(q, ref, kv_indptr, indices, last, "NHD")
    assert (o.float().cpu() - o_ref).abs().max().item() <= 1e-2
    assert (lse.cpu() - lse_ref).abs().max().item() <= 1e-3


@pytest.mark.gpu
def test_append_argument_errors():
    import flashinfer_b200 as fi

    cache = torch.zeros(4, 2, 16, 2, 64, dtype=torch.float16, device=DEV)
    k = torch.zeros(3, 2, 64, dtype=torch.float16, device=DEV)
    idx = torch.zeros(3, dtype=torch.int32, device=DEV)
    ind = torch.arange(4, dtype=torch.int32, device=DEV)
    ptr = torch.tensor([0, 4], dtype=torch.int32, device=DEV)
    last = torch.tensor([3], dtype=torch.int32, device=DEV)
    with pytest.raises(KeyError):
        fi.append_paged_kv_cache(k, k, idx, idx, cache, ind, ptr, last, "XYZ")
    with pytest.raises(ValueError):
        fi.append_paged_kv_cache(k, k.to(torch.bfloat16), idx, idx, cache, ind, ptr, last)
    with pytest.raises(ValueError):
        fi.append_paged_kv_cache(k[:, :1], k[:, :1], idx, idx, cache, ind, ptr, last)
    with pytest.raises(ValueError):
        fi.append_paged_kv_cache(k, k, idx[:2], idx, cache, ind, ptr, last)
    with pytest.raises(RuntimeError):
        fi.append_paged_kv_cache(k.cpu(), k.cpu(), idx.cpu(), idx.cpu(), cache.cpu(), ind.cpu(), ptr.cpu(), last.cpu())
