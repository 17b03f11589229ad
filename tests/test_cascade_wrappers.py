"""Cascade / shared-prefix wrappers (reference cascade.py:228-796,
tests/attention/test_shared_prefix_kernels.py): every level runs as per-row decode requests,
so the CPU part pins the row expansion and the GPU part compares the composed result with the
oracle's attention over the CONCATENATED KV of all levels (the identity cascade attention
rests on)."""
import math

import numpy as np
import pytest
import torch

from flashinfer_b200.cascade import rows_as_decode_requests

O_TOL = 1e-2  # north_star: max-abs 1e-2 for fp16 / bf16


def brute_rows(qo, kvp, kvi, lpl, page, causal):
    indptr, indices, last = [0], [], []
    for g in range(len(qo) - 1):
        n = qo[g + 1] - qo[g]
        pages = kvp[g + 1] - kvp[g]
        kv_len = (pages - 1) * page + lpl[g] if pages > 0 else 0
        for t in range(n):
            L = max(kv_len - (n - 1 - t), 0) if causal else kv_len
            np_ = -(-L // page)
            indices += list(kvi[kvp[g]:kvp[g] + np_])
            indptr.append(indptr[-1] + np_)
            last.append((L - 1) % page + 1 if L > 0 else 0)
    return indptr, indices, last


@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("page", [1, 5, 16])
def test_rows_as_decode_requests(causal, page):
    rng = np.random.default_rng(page + causal)
    groups = 6
    rows = rng.integers(0, 5, groups)
    rows[2] = 0
    pages = rng.integers(0, 7, groups)
    pages[0] = max(pages[0], 1)
    qo = np.concatenate([[0], np.cumsum(rows)])
    kvp = np.concatenate([[0], np.cumsum(pages)])
    kvi = rng.permutation(int(kvp[-1]) + 4)[: int(kvp[-1])]
    lpl = np.where(pages > 0, rng.integers(1, page + 1, groups), 0)
    got = rows_as_decode_requests(
        torch.tensor(qo, dtype=torch.int32), torch.tensor(kvp, dtype=torch.int32),
        torch.tensor(kvi, dtype=torch.int32), torch.tensor(lpl, dtype=torch.int32), page, causal)
    exp = brute_rows(qo, kvp, kvi, lpl, page, causal)
    assert got[0].dtype == got[1].dtype == got[2].dtype == torch.int32
    assert got[0].tolist() == exp[0]
    assert got[1].tolist() == [int(x) for x in exp[1]]
    assert got[2].tolist() == exp[2]


def test_rows_as_decode_requests_rejects_mismatch():
    i32 = lambda x: torch.tensor(x, dtype=torch.int32)
    with pytest.raises(ValueError):
        rows_as_decode_requests(i32([0, 1, 2]), i32([0, 1]), i32([0]), i32([1]), 16, False)
    with pytest.raises(ValueError):
        rows_as_decode_requests(i32([1, 2]), i32([0, 1]), i32([0]), i32([1]), 16, False)


# ------------------------------------------------------------------------------- GPU
def _full_attention(q_row, k, v, sm_scale):
    """fp32 attention of one query row [Hq, D] over [L, Hkv, D] (oracle: decode_attention_ref)."""
    from oracle.attention_ref import decode_attention_ref

    return decode_attention_ref(q_row.cpu(), k.float().cpu(), v.float().cpu(), sm_scale)[0]


@pytest.mark.gpu
@pytest.mark.parametrize("kv_layout", ["NHD", "HND"])
@pytest.mark.parametrize("shared_len,unique_len", [(128, 37), (2048, 17), (100, 1), (333, 64)])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_shared_prefix_decode_wrapper(kv_layout, shared_len, unique_len, dtype):
    """reference tests/attention/test_shared_prefix_kernels.py:40-170 (decode stage)."""
    import flashinfer_b200 as fi
    from helpers import make_paged_case, make_q

    dev = "cuda:0"
    bs, Hq, Hkv, D, page = 12, 32, 8, 128, 16
    lens = [unique_len + (i % 3) for i in range(bs)]
    cache, indptr, indices, last = make_paged_case(lens, Hkv, D, page, kv_layout, dtype, seed=3, device=dev)
    q = make_q(bs, Hq, D, dtype, seed=4, device=dev)
    g = torch.Generator().manual_seed(5)
    shape = (shared_len, Hkv, D) if kv_layout == "NHD" else (Hkv, shared_len, D)
    k_shared = (torch.randn(*shape, generator=g) / 2).to(dtype).to(dev)
    v_shared = torch.randn(*shape, generator=g).to(dtype).to(dev)
    ws = torch.zeros(32 << 20, dtype=torch.uint8, device=dev)
    w = fi.BatchDecodeWithSharedPrefixPagedKVCacheWrapper(ws, kv_layout)
    w.begin_forward(indptr.to(dev), indices.to(dev), last.to(dev), Hq, Hkv, D, page, data_type=dtype)
    o = w.forward(q, k_shared, v_shared, cache)
    o2 = w.forward(q, k_shared, v_shared, cache)  # cached shared-level plan
    assert torch.equal(o, o2)

    from oracle.attention_ref import gather_paged_kv, unpack_cache

    kc, vc = unpack_cache(cache.cpu(), kv_layout)
    ks = k_shared.cpu() if kv_layout == "NHD" else k_shared.cpu().permute(1, 0, 2)
    vs = v_shared.cpu() if kv_layout == "NHD" else v_shared.cpu().permute(1, 0, 2)
    for i in range(bs):
        ku, vu = gather_paged_kv(kc, vc, indptr.tolist(), indices.tolist(), last.tolist(), i, kv_layout)
        ref = _full_attention(q[i], torch.cat([ks, ku]), torch.cat([vs, vu]), 1.0 / math.sqrt(D))
        assert (o[i].float().cpu() - ref).abs().max().item() <= O_TOL


def _level_tables(page_lists, last_lens):
    indptr = torch.tensor([0] + list(np.cumsum([len(p) for p in page_lists])), dtype=torch.int32)
    indices = torch.tensor([x for p in page_lists for x in p], dtype=torch.int32)
    return indptr, indices, torch.tensor(last_lens, dtype=torch.int32)


@pytest.mark.gpu
@pytest.mark.parametrize("kv_layout", ["NHD", "HND"])
@pytest.mark.parametrize("causal,rows_per_req", [(False, 1), (True, 1), (True, 3)])
def test_multi_level_cascade(kv_layout, causal, rows_per_req):
    """Three levels over one page pool (reference test_shared_prefix_kernels.py:172-316 uses two;
    the docstring example cascade.py:243-296 is the two-level decode case): level 0 shared by
    every request, level 1 by halves of the batch, level 2 unique (causal append when
    rows_per_req > 1).  Compared with full attention over the concatenated KV per query row."""
    import flashinfer_b200 as fi

    dev = "cuda:0"
    dtype = torch.float16
    bs, Hq, Hkv, D, page = 6, 16, 4, 128, 16
    lens0, lens1 = [200], [75, 33]
    lens2 = [rows_per_req + 11 * i for i in range(bs)]  # includes the rows being appended
    need = lambda L: -(-L // page)
    total_pages = need(lens0[0]) + sum(need(x) for x in lens1) + sum(need(x) for x in lens2)
    g = torch.Generator().manual_seed(11)
    perm = torch.randperm(total_pages + 3, generator=g).tolist()
    take = lambda n: [perm.pop() for _ in range(n)]
    shape = (total_pages + 3, 2, page, Hkv, D) if kv_layout == "NHD" else (total_pages + 3, 2, Hkv, page, D)
    cache = torch.randn(*shape, generator=g).to(dtype).to(dev)
    lv_pages = [[take(need(L)) for L in lens] for lens in (lens0, lens1, lens2)]
    lv_last = [[(L - 1) % page + 1 for L in lens] for lens in (lens0, lens1, lens2)]
    tables = [_level_tables(p, l) for p, l in zip(lv_pages, lv_last)]
    n = rows_per_req
    qo = [torch.tensor([0, bs * n], dtype=torch.int32),
          torch.tensor([0, 3 * n, bs * n], dtype=torch.int32),
          torch.arange(bs + 1, dtype=torch.int32) * n]
    q = (torch.randn(bs * n, Hq, D, generator=g)).to(dtype).to(dev)

    ws = torch.zeros(32 << 20, dtype=torch.uint8, device=dev)
    w = fi.MultiLevelCascadeAttentionWrapper(3, ws, kv_layout)
    w.plan([x.to(dev) for x in qo], [t[0].to(dev) for t in tables], [t[1].to(dev) for t in tables],
           [t[2].to(dev) for t in tables], Hq, Hkv, D, page, causal=causal, q_data_type=dtype)
    o = w.run(q, cache)
    assert o.shape == q.shape and o.dtype == dtype

    from oracle.attention_ref import gather_paged_kv, unpack_cache

    kc, vc = unpack_cache(cache.cpu(), kv_layout)
    kv = [[gather_paged_kv(kc, vc, t[0].tolist(), t[1].tolist(), t[2].tolist(), i, kv_layout)
           for i in range(len(t[2]))] for t in tables]
    for r in range(bs):
        for t in range(n):
            row = r * n + t
            k2, v2 = kv[2][r]
            keep = lens2[r] - (n - 1 - t) if causal else lens2[r]
            k = torch.cat([kv[0][0][0], kv[1][0 if r < 3 else 1][0], k2[:keep]])
            v = torch.cat([kv[0][0][1], kv[1][0 if r < 3 else 1][1], v2[:keep]])
            ref = _full_attention(q[row], k, v, 1.0 / math.sqrt(D))
            assert (o[row].float().cpu() - ref).abs().max().item() <= O_TOL, (r, t)


@pytest.mark.gpu
def test_cascade_wrapper_contract():
    import flashinfer_b200 as fi

    ws = torch.zeros(8 << 20, dtype=torch.uint8, device="cuda:0")
    with pytest.raises(ValueError):  # graph mode needs the per-level page-capacity buffers
        fi.MultiLevelCascadeAttentionWrapper(2, ws, "NHD", use_cuda_graph=True)
    w = fi.MultiLevelCascadeAttentionWrapper(2, ws, "NHD")
    assert w.is_cuda_graph_enabled is False
    with pytest.raises(RuntimeError):
        w.run(torch.zeros(1, 8, 128, device="cuda:0", dtype=torch.float16), None)
    i32 = lambda x: torch.tensor(x, dtype=torch.int32)
    with pytest.raises(NotImplementedError):
        w.plan([i32([0, 1])] * 2, [i32([0, 1])] * 2, [i32([0])] * 2, [i32([1])] * 2, 8, 8, 128, 16,
               pos_encoding_mode="ROPE_LLAMA")
    with pytest.raises(ValueError):
        w.plan([i32([0, 1])], [i32([0, 1])], [i32([0])], [i32([1])], 8, 8, 128, 16)


@pytest.mark.gpu
@pytest.mark.parametrize("causal,rows_per_req", [(False, 1), (True, 2)])
def test_multi_level_cascade_in_a_cuda_graph(causal, rows_per_req):
    """use_cuda_graph=True (reference cascade.py:313-330): run() captured once, replayed after re-plans
    with other unique-level lengths; every replay against full attention over the concatenated KV."""
    import flashinfer_b200 as fi
    from oracle.attention_ref import gather_paged_kv, unpack_cache

    dev, dtype = "cuda:0", torch.float16
    bs, Hq, Hkv, D, page, n = 4, 16, 4, 128, 16, rows_per_req
    need = lambda L: -(-L // page)
    g = torch.Generator().manual_seed(21)
    pool = 200
    cache = torch.randn(pool, 2, page, Hkv, D, generator=g).to(dtype).to(dev)
    q = torch.randn(bs * n, Hq, D, generator=g).to(dtype).to(dev)
    ws = torch.zeros(32 << 20, dtype=torch.uint8, device=dev)
    cap = [torch.zeros(64, dtype=torch.int32, device=dev) for _ in range(2)]
    w = fi.MultiLevelCascadeAttentionWrapper(2, ws, "NHD", use_cuda_graph=True, paged_kv_indices_buf_arr=cap)
    assert w.is_cuda_graph_enabled
    qo = [torch.tensor([0, bs * n], dtype=torch.int32), torch.arange(bs + 1, dtype=torch.int32) * n]

    def tables(shared_len, lens, seed):
        perm = torch.randperm(pool, generator=torch.Generator().manual_seed(seed)).tolist()
        take = lambda k: [perm.pop() for _ in range(k)]
        lv = [[take(need(shared_len))], [take(need(L)) for L in lens]]
        last = [[(shared_len - 1) % page + 1], [(L - 1) % page + 1 for L in lens]]
        return [_level_tables(p, l) for p, l in zip(lv, last)]

    def check(o, tb, shared_len, lens):
        kc, vc = unpack_cache(cache.cpu(), "NHD")
        ks, vs = gather_paged_kv(kc, vc, tb[0][0].tolist(), tb[0][1].tolist(), tb[0][2].tolist(), 0, "NHD")
        for r in range(bs):
            k2, v2 = gather_paged_kv(kc, vc, tb[1][0].tolist(), tb[1][1].tolist(), tb[1][2].tolist(), r, "NHD")
            for t in range(n):
                keep = lens[r] - (n - 1 - t) if causal else lens[r]
                ref = _full_attention(q[r * n + t], torch.cat([ks, k2[:keep]]), torch.cat([vs, v2[:keep]]),
                                      1.0 / math.sqrt(D))
                assert (o[r * n + t].float().cpu() - ref).abs().max().item() <= O_TOL, (r, t)

    cases = [(100, [n + 5, 40, n, 130], 1), (333, [64, n + 1, 200, 17], 2), (16, [n, n, n + 30, 90], 3)]
    graph, out = None, None
    for shared_len, lens, seed in cases:
        tb = tables(shared_len, lens, seed)
        w.plan(qo, [t[0] for t in tb], [t[1] for t in tb], [t[2] for t in tb], Hq, Hkv, D, page, causal=causal,
               q_data_type=dtype)
        if graph is None:
            w.run(q, cache)  # warm-up outside the capture
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                out = w.run(q, cache)
        graph.replay()
        torch.cuda.synchronize()
        check(out, tb, shared_len, lens)
    with pytest.raises(ValueError):  # another number of query rows cannot reuse the captured shapes
        w.plan([torch.tensor([0, bs * n + 1], dtype=torch.int32), torch.arange(bs + 2, dtype=torch.int32)],
               [t[0] for t in tb], [t[1] for t in tb], [t[2] for t in tb], Hq, Hkv, D, page, q_data_type=dtype)


@pytest.mark.gpu
@pytest.mark.parametrize("kv_dtype", [torch.float8_e4m3fn, torch.bfloat16])
def test_two_level_cascade_shared_prefix_rows_share_the_stream(kv_dtype):
    """Shared prefix (one group, every request a row) + unique suffixes with 3 appended rows each: both levels
    run as multi-row units (fp8 KV: the tcgen05 kernel's wide mode, non-causal below the last level) —
    against full attention over the concatenated KV."""
    import flashinfer_b200 as fi
    from oracle.attention_ref import gather_paged_kv, unpack_cache

    dev, dtype = "cuda:0", torch.bfloat16
    bs, Hq, Hkv, D, page, n = 5, 16, 4, 128, 16, 3
    shared_len, lens = 777, [n + 40, n, 300, n + 1, 2000]
    need = lambda L: -(-L // page)
    g = torch.Generator().manual_seed(31)
    pool = need(shared_len) + sum(need(L) for L in lens) + 4
    cache = (torch.randn(pool, 2, page, Hkv, D, generator=g) * 0.5).to(kv_dtype).to(dev)
    q = torch.randn(bs * n, Hq, D, generator=g).to(dtype).to(dev)
    perm = torch.randperm(pool, generator=g).tolist()
    take = lambda k: [perm.pop() for _ in range(k)]
    tb = [_level_tables([take(need(shared_len))], [(shared_len - 1) % page + 1]),
          _level_tables([take(need(L)) for L in lens], [(L - 1) % page + 1 for L in lens])]
    qo = [torch.tensor([0, bs * n], dtype=torch.int32), torch.arange(bs + 1, dtype=torch.int32) * n]
    w = fi.MultiLevelCascadeAttentionWrapper(2, torch.zeros(32 << 20, dtype=torch.uint8, device=dev), "NHD")
    w.plan(qo, [t[0] for t in tb], [t[1] for t in tb], [t[2] for t in tb], Hq, Hkv, D, page, causal=True,
           q_data_type=dtype, kv_data_type=kv_dtype)
    assert [x._q_rows for x in w._batch_decode_wrappers] == [bs * n, n]
    assert [x._q_rows_noncausal for x in w._batch_decode_wrappers] == [1, 0]
    o = w.run(q, cache)
    kc, vc = unpack_cache(cache.cpu(), "NHD")
    ks, vs = gather_paged_kv(kc, vc, tb[0][0].tolist(), tb[0][1].tolist(), tb[0][2].tolist(), 0, "NHD")
    tol = 3e-2 if kv_dtype != torch.bfloat16 else O_TOL
    for r in range(bs):
        k2, v2 = gather_paged_kv(kc, vc, tb[1][0].tolist(), tb[1][1].tolist(), tb[1][2].tolist(), r, "NHD")
        for t in range(n):
            keep = lens[r] - (n - 1 - t)
            ref = _full_attention(q[r * n + t], torch.cat([ks, k2[:keep]]).float(), torch.cat([vs, v2[:keep]]).float(),
                                  1.0 / math.sqrt(D))
            assert (o[r * n + t].float().cpu() - ref).abs().max().item() <= tol, (r, t)
