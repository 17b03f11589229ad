"""GPU parity tests of the decode path — through the public API, i.e. through the C-ABI.

Oracle: oracle/attention_ref.py (fp32 CPU restatement pinned to the reference's test oracle).
Gates (BASELINE.json north_star): o max-abs <= 1e-2 (fp16/bf16 KV), <= 3e-2 (fp8 KV);
base-2 lse within 1e-3.  Cases follow the reference's matrix
(tests/attention/test_batch_decode_kernels.py:58-188, test_decode_prefill_lse.py:22-71,
test_non_contiguous_decode.py:46-118, test_decode_fp8_calibration_scale.py:74-153)."""
import math

import numpy as np
import pytest
import torch

from helpers import golden_cases, golden_inputs, make_paged_case, make_q
from oracle import attention_ref as orc

pytestmark = pytest.mark.gpu

O_TOL = 1e-2
O_TOL_FP8 = 3e-2
LSE_TOL = 1e-3
DEV = "cuda:0"


def _fi():
    import flashinfer_b200 as fi

    return fi


def _workspace(nbytes=128 << 20):
    return torch.zeros(nbytes, dtype=torch.uint8, device=DEV)


def run_batch(kv_lens, hq, hkv, d, page, layout, q_dtype, kv_dtype=None, seed=0, tuple_cache=False,
              loader=0, window_left=-1, soft_cap=None, ws=None, check_requests=None):
    fi = _fi()
    from flashinfer_b200 import _lib

    kv_dtype = kv_dtype or q_dtype
    is8 = kv_dtype in (torch.float8_e4m3fn, torch.float8_e5m2)
    cache, indptr, indices, last = make_paged_case(kv_lens, hkv, d, page, layout,
                                                   torch.float32, seed=seed)
    k_scale = v_scale = None
    if is8:
        # test_decode_fp8_calibration_scale.py:131-137: scale = amax / 256
        k_scale = float(cache[:, 0].abs().amax() / 256)
        v_scale = float(cache[:, 1].abs().amax() / 256)
        cache[:, 0] /= k_scale
        cache[:, 1] /= v_scale
    cache = cache.to(kv_dtype)
    q = make_q(len(kv_lens), hq, d, q_dtype, seed=seed + 1)
    cache_d, q_d = cache.to(DEV), q.to(DEV)
    wrapper = fi.BatchDecodeWithPagedKVCacheWrapper(ws if ws is not None else _workspace(), layout)
    _lib.load().fi_set_tuning(0, 0, loader, 0)
    try:
        wrapper.plan(indptr, indices, last, hq, hkv, d, page, q_data_type=q_dtype,
                     kv_data_type=kv_dtype, window_left=window_left, logits_soft_cap=soft_cap)
        pkv = (cache_d[:, 0], cache_d[:, 1]) if tuple_cache else cache_d
        o, lse = wrapper.run(q_d, pkv, k_scale=k_scale, v_scale=v_scale, return_lse=True)
        torch.cuda.synchronize()
    finally:
        _lib.load().fi_set_tuning(0, 0, 0, 0)
    o_ref, lse_ref = orc.batch_decode_ref(
        q, cache, indptr, indices, last, layout, k_scale=k_scale or 1.0, v_scale=v_scale or 1.0,
        logits_soft_cap=soft_cap or 0.0, window_left=window_left, requests=check_requests)
    o_c, lse_c = o.float().cpu(), lse.cpu()
    if check_requests is not None:
        o_c, lse_c = o_c[list(check_requests)], lse_c[list(check_requests)]
    tol = O_TOL_FP8 if is8 else O_TOL
    assert (o_c - o_ref).abs().max().item() <= tol, (o_c - o_ref).abs().max().item()
    assert (lse_c - lse_ref).abs().max().item() <= LSE_TOL, (lse_c - lse_ref).abs().max().item()
    return o, lse


@pytest.mark.parametrize("loader", [1, 2])  # 1 = TMA boxes, 2 = cp.async gathers
@pytest.mark.parametrize("layout", ["NHD", "HND"])
@pytest.mark.parametrize("q_dtype", [torch.float16, torch.bfloat16])
def test_batch_decode_minimum_slice(loader, layout, q_dtype):
    # bs in {1, 12, 17}, kv_len in {1, 54, 97, 512, 2048}, shuffled pages
    run_batch([1, 54, 97, 512, 2048, 33, 64, 65, 128, 1000, 17, 3], 32, 8, 128, 16, layout,
              q_dtype, loader=loader)


@pytest.mark.parametrize("page", [1, 5, 8, 16, 32])
@pytest.mark.parametrize("group", [1, 2, 4, 8])
def test_batch_decode_page_sizes_and_groups(page, group):
    hkv = 4
    run_batch([54, 97, 512, 1, 2048, 130], hkv * group, hkv, 128, page, "NHD", torch.float16,
              seed=page * 10 + group)


@pytest.mark.parametrize("hq,hkv", [(16, 1), (32, 2), (24, 2), (64, 2), (9, 1), (40, 4)])
@pytest.mark.parametrize("kv_dtype", [None, torch.float8_e4m3fn])
def test_batch_decode_large_gqa_groups(hq, hkv, kv_dtype):
    """GQA groups above one MMA N tile (16, 12, 32, 9, 10) run as virtual KV heads of <= 8 query
    heads that address the same cache head (the reference's CUDA-core decode stops at group 8,
    utils.cuh:123-143; its tensor-core decode, which engines pick for large groups, takes any)."""
    fi = _fi()
    lens = [1, 97, 512, 3000, 0, 64]
    d, page = 128, 16
    is8 = kv_dtype is not None
    cache, indptr, indices, last = make_paged_case(lens, hkv, d, page, "NHD", torch.float32, seed=61)
    cache = (cache * 0.5 if is8 else cache).to(kv_dtype or torch.float16)
    q = make_q(len(lens), hq, d, torch.float16, seed=62)
    w = fi.BatchDecodeWithPagedKVCacheWrapper(_workspace(), "NHD", use_tensor_cores=True)
    w.plan(indptr, indices, last, hq, hkv, d, page, q_data_type=torch.float16, kv_data_type=cache.dtype)
    o, lse = w.run(q.to(DEV), cache.to(DEV), return_lse=True)
    torch.cuda.synchronize()
    o_ref, lse_ref = orc.batch_decode_ref(q, cache, indptr, indices, last, "NHD")
    assert (o.float().cpu() - o_ref).abs().max().item() <= (O_TOL_FP8 if is8 else O_TOL)
    assert (lse.cpu() - lse_ref).abs().max().item() <= LSE_TOL
    # single decode takes the same route
    k_i, v_i = orc.gather_paged_kv(*orc.unpack_cache(cache, "NHD"), indptr, indices, last, 3, "NHD")
    o_i = fi.single_decode_with_kv_cache(q[3].to(DEV), k_i.to(DEV), v_i.to(DEV))
    assert (o_i.float().cpu() - o_ref[3]).abs().max().item() <= (O_TOL_FP8 if is8 else O_TOL)


@pytest.mark.parametrize("d", [64, 128, 256])
@pytest.mark.parametrize("layout", ["NHD", "HND"])
def test_batch_decode_head_dims(d, layout):
    run_batch([97, 512, 54, 2047], 8, 2, d, 8, layout, torch.bfloat16, seed=d)


@pytest.mark.parametrize("bs,kv_len", [(12, 54), (17, 97), (128, 512), (12, 2048), (3, 16384)])
def test_batch_decode_reference_matrix(bs, kv_len):
    # test_batch_decode_kernels.py:58-66 (uniform lengths)
    run_batch([kv_len] * bs, 32, 4, 128, 16, "NHD", torch.float16, seed=bs,
              check_requests=range(min(bs, 6)))


def test_batch_decode_tuple_cache_and_user_buffers():
    fi = _fi()
    kv_lens = [100, 7, 300]
    run_batch(kv_lens, 8, 2, 128, 16, "HND", torch.float16, tuple_cache=True)
    cache, indptr, indices, last = make_paged_case(kv_lens, 2, 128, 16, "NHD", torch.float16)
    q = make_q(3, 8, 128, torch.float16).to(DEV)
    w = fi.BatchDecodeWithPagedKVCacheWrapper(_workspace(), "NHD")
    w.plan(indptr, indices, last, 8, 2, 128, 16, q_data_type=torch.float16)
    o = w.run(q, cache.to(DEV))
    out = torch.empty_like(q)
    lse = torch.empty(3, 8, dtype=torch.float32, device=DEV)
    o2, lse2 = w.run(q, cache.to(DEV), out=out, lse=lse, return_lse=True)
    assert o2.data_ptr() == out.data_ptr() and torch.equal(o, out)
    with pytest.raises(ValueError):
        w.run(q.to(torch.bfloat16), cache.to(DEV))
    with pytest.raises(ValueError):
        w.run(q, cache.to(DEV), out=torch.empty(3, 8, 64, dtype=torch.float16, device=DEV))
    with pytest.raises(ValueError):
        w.plan(indptr.long(), indices, last, 8, 2, 128, 16)


def test_empty_request_regression():
    # tests/attention/test_decode_prefill_lse.py:24-26: indptr=[0,0,9], last_page_len=[0,1], HND
    fi = _fi()
    hq, hkv, d, page = 8, 2, 128, 16
    g = torch.Generator().manual_seed(0)
    cache = torch.randn(9, 2, hkv, page, d, generator=g).half()
    indptr = torch.tensor([0, 0, 9], dtype=torch.int32)
    indices = torch.arange(9, dtype=torch.int32)
    last = torch.tensor([0, 1], dtype=torch.int32)
    q = make_q(2, hq, d, torch.float16)
    w = fi.BatchDecodeWithPagedKVCacheWrapper(_workspace(), "HND")
    w.plan(indptr, indices, last, hq, hkv, d, page, q_data_type=torch.float16)
    o, lse = w.run(q.to(DEV), cache.to(DEV), return_lse=True)
    o_ref, lse_ref = orc.batch_decode_ref(q, cache, indptr, indices, last, "HND")
    assert torch.all(o[0] == 0) and torch.all(lse[0] == -5e4)
    assert (o.float().cpu() - o_ref).abs().max() <= O_TOL
    assert (lse.cpu() - lse_ref).abs().max() <= LSE_TOL


@pytest.mark.parametrize("kv_dtype", [torch.float8_e4m3fn, torch.float8_e5m2])
@pytest.mark.parametrize("layout", ["NHD", "HND"])
@pytest.mark.parametrize("loader", [1, 2])
def test_batch_decode_fp8_kv(kv_dtype, layout, loader):
    run_batch([54, 97, 512, 2048, 1, 77], 32, 4, 128, 16, layout, torch.float16, kv_dtype,
              loader=loader)
    run_batch([300, 5000], 8, 1, 128, 16, layout, torch.bfloat16, kv_dtype, loader=loader)


def test_non_contiguous_q_and_cache():
    # test_non_contiguous_decode.py:100-118: q sliced out of a packed QKV projection
    fi = _fi()
    hq, hkv, d, page = 8, 2, 128, 16
    kv_lens = [70, 200, 5]
    cache, indptr, indices, last = make_paged_case(kv_lens, hkv, d, page, "NHD", torch.float16)
    g = torch.Generator().manual_seed(5)
    qkv = torch.randn(3, (hq + 2 * hkv) * d, generator=g).half().to(DEV)
    q = qkv[:, : hq * d].view(3, hq, d)
    assert not q.is_contiguous()
    # cache rows padded: strided view on the head_dim-major storage
    big = torch.zeros(cache.shape[0], 2, page, hkv, 2 * d, dtype=torch.float16, device=DEV)
    big[..., :d] = cache.to(DEV)
    view = big[..., :d]
    w = fi.BatchDecodeWithPagedKVCacheWrapper(_workspace(), "NHD")
    w.plan(indptr, indices, last, hq, hkv, d, page, q_data_type=torch.float16)
    o, lse = w.run(q, view, return_lse=True)
    o_ref, lse_ref = orc.batch_decode_ref(q.cpu(), cache, indptr, indices, last, "NHD")
    assert (o.float().cpu() - o_ref).abs().max() <= O_TOL
    assert (lse.cpu() - lse_ref).abs().max() <= LSE_TOL


def test_window_left_and_soft_cap():
    run_batch([300, 10, 1000, 64], 8, 2, 128, 16, "NHD", torch.float16, window_left=99)
    run_batch([300, 10, 1000, 64], 8, 2, 128, 16, "NHD", torch.float16, soft_cap=30.0)


def test_split_kv_long_single_request_and_determinism():
    # one long request: every CTA gets a slice, one merge unit per head
    o1, l1 = run_batch([50000], 32, 8, 128, 16, "NHD", torch.bfloat16)
    o2, l2 = run_batch([50000], 32, 8, 128, 16, "NHD", torch.bfloat16)
    assert torch.equal(o1, o2) and torch.equal(l1, l2)  # fixed inputs + plan => bit-identical


def test_cuda_graph_replay_with_replan():
    # test_batch_decode_kernels.py:492-676: capture run, re-plan between replays
    fi = _fi()
    hq, hkv, d, page, bs = 32, 8, 128, 16, 4
    max_pages = 4096
    indptr_buf = torch.zeros(bs + 1, dtype=torch.int32, device=DEV)
    indices_buf = torch.zeros(max_pages, dtype=torch.int32, device=DEV)
    last_buf = torch.zeros(bs, dtype=torch.int32, device=DEV)
    w = fi.CUDAGraphBatchDecodeWithPagedKVCacheWrapper(_workspace(), indptr_buf, indices_buf,
                                                       last_buf, "NHD")
    lens_a, lens_b = [100, 2000, 31, 977], [4000, 1, 64, 15000]
    cache, ip_b, ix_b, lp_b = make_paged_case(lens_b, hkv, d, page, "NHD", torch.float16, seed=2)
    cache_d = cache.to(DEV)
    # plan A uses a page table that is valid inside the same pool
    _, ip_a, ix_a, lp_a = make_paged_case(lens_a, hkv, d, page, "NHD", torch.float16, seed=3)
    q = make_q(bs, hq, d, torch.float16).to(DEV)
    out = torch.empty_like(q)
    w.plan(ip_a, ix_a, lp_a, hq, hkv, d, page, q_data_type=torch.float16)
    w.run(q, cache_d, out=out)  # warm-up outside the graph
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        w.run(q, cache_d, out=out)
    for ip, ix, lp in ((ip_a, ix_a, lp_a), (ip_b, ix_b, lp_b), (ip_a, ix_a, lp_a)):
        w.plan(ip, ix, lp, hq, hkv, d, page, q_data_type=torch.float16)
        graph.replay()
        torch.cuda.synchronize()
        o_ref, _ = orc.batch_decode_ref(q.cpu(), cache, ip, ix, lp, "NHD")
        assert (out.float().cpu() - o_ref).abs().max() <= O_TOL


Z, NAMES = golden_cases()


@pytest.mark.parametrize("name", NAMES)
def test_single_decode_against_reference_golden(name):
    fi = _fi()
    q, k, v, ok = golden_inputs(Z, name)
    if not ok:
        pytest.skip("un-stored golden inputs not reproducible on this machine")
    o, lse = fi.single_decode_with_kv_cache(q.to(DEV), k.to(DEV), v.to(DEV), return_lse=True)
    assert np.abs(o.float().cpu().numpy() - Z[f"{name}/o_fp32"]).max() <= O_TOL
    assert np.abs(lse.cpu().numpy() - Z[f"{name}/lse2"]).max() <= LSE_TOL
    # HND layout gives the same answer
    o2 = fi.single_decode_with_kv_cache(q.to(DEV), k.transpose(0, 1).contiguous().to(DEV),
                                        v.transpose(0, 1).contiguous().to(DEV), kv_layout="HND")
    assert (o2.float() - o.float()).abs().max() <= 2e-3


@pytest.mark.parametrize("kv_len", [0, 1, 63, 64, 65, 255, 256, 257, 4099, 33001])
def test_single_decode_lengths(kv_len):
    fi = _fi()
    g = torch.Generator().manual_seed(kv_len)
    q = torch.randn(32, 128, generator=g).half()
    k = torch.randn(kv_len, 8, 128, generator=g).half()
    v = torch.randn(kv_len, 8, 128, generator=g).half()
    o, lse = fi.single_decode_with_kv_cache(q.to(DEV), k.to(DEV), v.to(DEV), return_lse=True)
    o_ref, lse_ref = orc.decode_attention_ref(q, k, v)
    assert (o.float().cpu() - o_ref).abs().max() <= O_TOL
    assert (lse.cpu() - lse_ref).abs().max() <= LSE_TOL


def test_batch_equals_single():
    # the reference's own chain: batch paged decode == single decode per request, 1e-3
    fi = _fi()
    kv_lens = [54, 512, 2048]
    cache, indptr, indices, last = make_paged_case(kv_lens, 4, 128, 16, "NHD", torch.float16)
    q = make_q(3, 32, 128, torch.float16)
    w = fi.BatchDecodeWithPagedKVCacheWrapper(_workspace(), "NHD")
    w.plan(indptr, indices, last, 32, 4, 128, 16, q_data_type=torch.float16)
    o = w.run(q.to(DEV), cache.to(DEV))
    for i in range(3):
        k, v = orc.gather_paged_kv(cache[:, 0], cache[:, 1], indptr, indices, last, i, "NHD")
        oi = fi.single_decode_with_kv_cache(q[i].to(DEV), k.contiguous().to(DEV), v.contiguous().to(DEV))
        torch.testing.assert_close(o[i], oi, rtol=1e-3, atol=1e-3)


@pytest.mark.parametrize("kv_dtype", [None, torch.float8_e4m3fn])
@pytest.mark.parametrize("loader", [1, 2])
def test_many_short_requests_refill_the_segment_queue(kv_dtype, loader):
    """Hundreds of one-tile (and empty) requests: work items hold far more segments than the three
    inline ones and than the 32-entry shared-memory queue, so the feeder streams them from the list."""
    rng = np.random.default_rng(7)
    kv_lens = rng.integers(0, 131, 700).tolist()
    kv_lens[5:45] = [0] * 40  # a run of empty requests longer than the queue
    run_batch(kv_lens, 8, 2, 128, 16, "NHD", torch.bfloat16, kv_dtype=kv_dtype, loader=loader,
              check_requests=list(range(0, 700, 9)))


@pytest.mark.parametrize("knobs", ["70,50,2,2", "90,50,4,4", "50,25,1,0"])
def test_dynamic_item_scheduler(knobs):
    """The guided self-scheduling mode (items drawn from the ticket counter, segment queue, deferred
    merge arrivals) is off by default and its knobs are read once per process: run the batch
    parity tests again in a child process with the scheduler switched on."""
    import os
    import subprocess
    import sys

    if os.environ.get("FI_B200_TEST_CHILD") == "1":
        pytest.skip("already inside the child run")
    f0, g, tail, look = knobs.split(",")
    env = dict(os.environ, FI_B200_TEST_CHILD="1", FI_B200_SCHED_F0=f0, FI_B200_SCHED_G=g,
               FI_B200_SCHED_TAIL=tail, FI_B200_SCHED_LOOKAHEAD=look)
    sel = ("test_batch_decode_minimum_slice or test_batch_decode_reference_matrix or "
           "test_empty_request_regression or test_batch_decode_fp8_kv or "
           "test_split_kv_long_single_request_and_determinism or test_cuda_graph_replay_with_replan "
           "or test_batch_decode_page_sizes_and_groups or test_many_short_requests")
    r = subprocess.run([sys.executable, "-m", "pytest", __file__, "-x", "-q", "-m", "gpu", "-k", sel],
                       env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert " passed" in r.stdout


@pytest.mark.parametrize("kv_dtype", [torch.bfloat16, torch.float8_e4m3fn])
def test_fixed_split_size_outputs_are_batch_invariant(kv_dtype):
    """fixed_split_size (reference decode.py:880-885): the same request gives bit-identical o / lse
    whether it is decoded alone or inside a batch of other requests, and matches the oracle."""
    fi = _fi()
    hq, hkv, d, page, split_pages = 32, 8, 128, 16, 24   # 384-token chunks (not tile-aligned)
    lens = [77, 9000, 5000, 1, 0, 2048, 385, 384]
    target = 2
    cache, indptr, indices, last = make_paged_case(lens, hkv, d, page, "NHD", torch.float32, seed=21)
    cache = cache.to(kv_dtype)
    q = make_q(len(lens), hq, d, torch.bfloat16, seed=22)
    cache_d, q_d = cache.to(DEV), q.to(DEV)
    w = fi.BatchDecodeWithPagedKVCacheWrapper(_workspace(), "NHD", use_tensor_cores=True)
    kw = dict(q_data_type=torch.bfloat16, kv_data_type=kv_dtype, fixed_split_size=split_pages)
    w.plan(indptr, indices, last, hq, hkv, d, page, **kw)
    o_b, lse_b = w.run(q_d, cache_d, return_lse=True)
    # the target request alone: same pages, a one-request page table
    lo, hi = int(indptr[target]), int(indptr[target + 1])
    w.plan(torch.tensor([0, hi - lo], dtype=torch.int32), indices[lo:hi].contiguous(),
           last[target:target + 1].contiguous(), hq, hkv, d, page, **kw)
    o_a, lse_a = w.run(q_d[target:target + 1], cache_d, return_lse=True)
    torch.cuda.synchronize()
    assert torch.equal(o_a[0], o_b[target]) and torch.equal(lse_a[0], lse_b[target])
    o_ref, lse_ref = orc.batch_decode_ref(q, cache, indptr, indices, last, "NHD")
    tol = O_TOL_FP8 if kv_dtype != torch.bfloat16 else O_TOL
    assert (o_b.float().cpu() - o_ref).abs().max().item() <= tol
    assert (lse_b.cpu() - lse_ref).abs().max().item() <= LSE_TOL
    # not a tensor-core wrapper: same error as the reference (decode.py:966-969)
    w2 = fi.BatchDecodeWithPagedKVCacheWrapper(_workspace(), "NHD")
    with pytest.raises(ValueError):
        w2.plan(indptr, indices, last, hq, hkv, d, page, **kw)


def test_disable_split_kv_matches_oracle_and_never_merges():
    fi = _fi()
    hq, hkv, d, page = 16, 4, 128, 16
    lens = [20000, 3, 0, 900, 64]
    cache, indptr, indices, last = make_paged_case(lens, hkv, d, page, "HND", torch.float16, seed=31)
    q = make_q(len(lens), hq, d, torch.float16, seed=32)
    w = fi.BatchDecodeWithPagedKVCacheWrapper(_workspace(), "HND")
    w.plan(indptr, indices, last, hq, hkv, d, page, q_data_type=torch.float16, disable_split_kv=True)
    assert int(w._plan_info[3]) == 0 and int(w._plan_info[4]) == 0   # no merge units, no partial slots
    o, lse = w.run(q.to(DEV), cache.to(DEV), return_lse=True)
    o_ref, lse_ref = orc.batch_decode_ref(q, cache, indptr, indices, last, "HND")
    assert (o.float().cpu() - o_ref).abs().max().item() <= O_TOL
    assert (lse.cpu() - lse_ref).abs().max().item() <= LSE_TOL


# ---------------------------------------------------------------- ALiBi
def _alibi_golden():
    from pathlib import Path

    z = np.load(Path(__file__).resolve().parent / "golden" / "alibi_golden.npz")
    return z, [str(n) for n in z["names"]]


def _alibi_lse_ok(lse, lse_ref):
    # ALiBi logits grow with the position (slope * 4098 * log2 e ~ 5000 in base 2), where one f32
    # ulp is already 5e-4: the 1e-3 gate gets a 4-ulp relative term
    return bool(((lse - lse_ref).abs() <= LSE_TOL + lse_ref.abs() * 2.0 ** -21).all())


@pytest.mark.parametrize("name", _alibi_golden()[1])
def test_single_decode_alibi_against_reference_golden(name):
    """reference tests/attention/test_alibi.py:55-70 with its oracle's outputs as fixtures"""
    fi = _fi()
    z, _ = _alibi_golden()
    H, D, L, is_bf16 = [int(x) for x in z[f"{name}/meta"]]
    dt = torch.bfloat16 if is_bf16 else torch.float16
    q, k, v, o_gold = (torch.from_numpy(z[f"{name}/{key}"].copy()).view(dt) for key in ("q", "k", "v", "o"))
    o = fi.single_decode_with_kv_cache(q.to(DEV), k.to(DEV), v.to(DEV), pos_encoding_mode="ALIBI")
    torch.testing.assert_close(o.cpu().float(), o_gold.float(), rtol=1e-3 if not is_bf16 else 1e-2,
                               atol=1e-3 if not is_bf16 else 1e-2)
    o2, lse = fi.single_decode_with_kv_cache(q.to(DEV), k.to(DEV), v.to(DEV), pos_encoding_mode="ALIBI",
                                             return_lse=True)
    slopes = torch.from_numpy(z[f"{name}/slopes"].copy())
    o_ref, lse_ref = orc.decode_attention_ref(q, k, v, alibi_slopes=slopes)
    assert (o2.float().cpu() - o_ref).abs().max().item() <= O_TOL
    assert _alibi_lse_ok(lse.cpu(), lse_ref)


@pytest.mark.parametrize("kv_dtype,loader", [(None, 1), (None, 2), (torch.float8_e4m3fn, 1),
                                             (torch.float8_e4m3fn, 2)])
@pytest.mark.parametrize("soft_cap,window_left", [(None, -1), (30.0, -1), (None, 300)])
def test_batch_decode_alibi(kv_dtype, loader, soft_cap, window_left):
    fi = _fi()
    from flashinfer_b200 import _lib
    from flashinfer_b200.utils import get_alibi_slopes

    hq, hkv, d, page = 32, 8, 128, 16
    lens = [1, 54, 97, 512, 2048, 0, 4099, 777]
    q_dtype = torch.bfloat16
    kvd = kv_dtype or q_dtype
    is8 = kv_dtype is not None
    cache, indptr, indices, last = make_paged_case(lens, hkv, d, page, "NHD", torch.float32, seed=41)
    k_scale = v_scale = None
    if is8:
        k_scale = float(cache[:, 0].abs().amax() / 256)
        v_scale = float(cache[:, 1].abs().amax() / 256)
        cache[:, 0] /= k_scale
        cache[:, 1] /= v_scale
    cache = cache.to(kvd)
    q = make_q(len(lens), hq, d, q_dtype, seed=42)
    w = fi.BatchDecodeWithPagedKVCacheWrapper(_workspace(), "NHD")
    _lib.load().fi_set_tuning(0, 0, loader, 0)
    try:
        w.plan(indptr, indices, last, hq, hkv, d, page, pos_encoding_mode="ALIBI", q_data_type=q_dtype,
               kv_data_type=kvd, logits_soft_cap=soft_cap, window_left=window_left)
        o, lse = w.run(q.to(DEV), cache.to(DEV), k_scale=k_scale, v_scale=v_scale, return_lse=True)
        torch.cuda.synchronize()
    finally:
        _lib.load().fi_set_tuning(0, 0, 0, 0)
    o_ref, lse_ref = orc.batch_decode_ref(q, cache, indptr, indices, last, "NHD", k_scale=k_scale or 1.0,
                                          v_scale=v_scale or 1.0, logits_soft_cap=soft_cap or 0.0,
                                          window_left=window_left, alibi_slopes=get_alibi_slopes(hq))
    assert (o.float().cpu() - o_ref).abs().max().item() <= (O_TOL_FP8 if is8 else O_TOL)
    assert _alibi_lse_ok(lse.cpu(), lse_ref), (lse.cpu() - lse_ref).abs().max().item()
    # the mode given to plan() sticks; NONE afterwards gives a different answer again
    w.plan(indptr, indices, last, hq, hkv, d, page, q_data_type=q_dtype, kv_data_type=kvd,
           logits_soft_cap=soft_cap, window_left=window_left)
    o_none = w.run(q.to(DEV), cache.to(DEV), k_scale=k_scale, v_scale=v_scale)
    assert (o_none.float().cpu() - o_ref).abs().max().item() > 1e-2


# ---------------------------------------------------------------- RoPE on the fly
def _rope_golden():
    from pathlib import Path

    z = np.load(Path(__file__).resolve().parent / "golden" / "rope_golden.npz")
    return z, [str(n) for n in z["names"]]


@pytest.mark.parametrize("name", _rope_golden()[1])
def test_single_decode_rope_against_reference_golden(name):
    fi = _fi()
    z, _ = _rope_golden()
    Hq, Hkv, D, L, is_bf16 = [int(x) for x in z[f"{name}/meta"]]
    dt = torch.bfloat16 if is_bf16 else torch.float16
    q, k, v = (torch.from_numpy(z[f"{name}/{key}"].copy()).view(dt) for key in ("q", "k", "v"))
    theta = float(z[f"{name}/theta"][0])
    o, lse = fi.single_decode_with_kv_cache(q.to(DEV), k.to(DEV), v.to(DEV), pos_encoding_mode="ROPE_LLAMA",
                                            rope_theta=theta, return_lse=True)
    assert (o.float().cpu() - torch.from_numpy(z[f"{name}/o"].copy())).abs().max().item() <= O_TOL
    assert (lse.cpu() - torch.from_numpy(z[f"{name}/lse2"].copy())).abs().max().item() <= LSE_TOL
    # HND layout gives the same answer
    o2 = fi.single_decode_with_kv_cache(q.to(DEV), k.transpose(0, 1).contiguous().to(DEV),
                                        v.transpose(0, 1).contiguous().to(DEV), kv_layout="HND",
                                        pos_encoding_mode="ROPE_LLAMA", rope_theta=theta)
    assert torch.equal(o2, o)


@pytest.mark.parametrize("layout,kv_dtype,page,d", [("NHD", None, 16, 128), ("HND", None, 8, 128),
                                                     ("NHD", torch.float8_e4m3fn, 16, 128),
                                                     ("NHD", None, 5, 64), ("HND", None, 16, 256)])
@pytest.mark.parametrize("soft_cap,window_left", [(None, -1), (30.0, 300)])
def test_batch_decode_rope(layout, kv_dtype, page, d, soft_cap, window_left):
    """reference tests/attention/test_batch_decode_kernels.py:58-185 (pos_encoding_mode="ROPE_LLAMA"):
    long requests are split over CTAs and merged, short and empty ones are not"""
    fi = _fi()
    hq, hkv = 16, 4
    lens = [1, 54, 97, 512, 2048, 0, 9000, 777]
    q_dtype = torch.float16
    kvd = kv_dtype or q_dtype
    is8 = kv_dtype is not None
    cache, indptr, indices, last = make_paged_case(lens, hkv, d, page, layout, torch.float32, seed=51)
    k_scale = v_scale = None
    if is8:
        k_scale = float(cache[:, 0].abs().amax() / 256)
        v_scale = float(cache[:, 1].abs().amax() / 256)
        cache[:, 0] /= k_scale
        cache[:, 1] /= v_scale
    cache = cache.to(kvd)
    q = make_q(len(lens), hq, d, q_dtype, seed=52)
    w = fi.BatchDecodeWithPagedKVCacheWrapper(_workspace(), layout)
    w.plan(indptr, indices, last, hq, hkv, d, page, pos_encoding_mode="ROPE_LLAMA", q_data_type=q_dtype,
           kv_data_type=kvd, logits_soft_cap=soft_cap, window_left=window_left, rope_theta=1e4)
    assert int(w._plan_info[3]) > 0   # some unit is split: the fused merge runs
    o, lse = w.run(q.to(DEV), cache.to(DEV), k_scale=k_scale, v_scale=v_scale, return_lse=True)
    torch.cuda.synchronize()
    o_ref, lse_ref = orc.batch_decode_ref(q, cache, indptr, indices, last, layout, k_scale=k_scale or 1.0,
                                          v_scale=v_scale or 1.0, logits_soft_cap=soft_cap or 0.0,
                                          window_left=window_left, rope=(1.0, 1e4))
    assert (o.float().cpu() - o_ref).abs().max().item() <= (O_TOL_FP8 if is8 else O_TOL)
    assert (lse.cpu() - lse_ref).abs().max().item() <= LSE_TOL
    # batch == single (the reference test's own check, test_batch_decode_kernels.py:176-183)
    if not is8 and soft_cap is None and window_left < 0:
        for i in (2, 4):
            k_i, v_i = orc.gather_paged_kv(*orc.unpack_cache(cache, layout), indptr, indices, last, i, layout)
            o_i = fi.single_decode_with_kv_cache(q[i].to(DEV), k_i.to(DEV), v_i.to(DEV),
                                                 pos_encoding_mode="ROPE_LLAMA")
            torch.testing.assert_close(o[i], o_i, rtol=1e-3, atol=1e-3)


def test_invalid_pos_encoding_mode():
    fi = _fi()
    q = torch.zeros(8, 128, dtype=torch.float16, device=DEV)
    k = torch.zeros(16, 2, 128, dtype=torch.float16, device=DEV)
    with pytest.raises(KeyError):
        fi.single_decode_with_kv_cache(q, k, k, pos_encoding_mode="SOMETHING")


@pytest.mark.parametrize("q_len_per_req", [2, 4])
@pytest.mark.parametrize("kv_dtype,pos_mode", [(None, "NONE"), (torch.float8_e4m3fn, "NONE"), (None, "ROPE_LLAMA")])
def test_multi_token_decode_q_len_per_req(q_len_per_req, kv_dtype, pos_mode):
    """run(q_len_per_req = n) (reference decode.py:1177,1278): n query tokens per
    request whose K / V are already in the cache; token t attends the first kv_len - (n - 1 - t) tokens."""
    fi = _fi()
    hq, hkv, d, page = 16, 4, 128, 16
    lens = [5, 97, 512, 3000, 64, 4]
    n = q_len_per_req
    is8 = kv_dtype is not None
    cache, indptr, indices, last = make_paged_case(lens, hkv, d, page, "NHD", torch.float32, seed=71)
    cache = (cache * 0.5 if is8 else cache).to(kv_dtype or torch.float16)
    q = make_q(len(lens) * n, hq, d, torch.float16, seed=72)
    w = fi.BatchDecodeWithPagedKVCacheWrapper(_workspace(), "NHD")
    w.plan(indptr, indices, last, hq, hkv, d, page, pos_encoding_mode=pos_mode, q_data_type=torch.float16,
           kv_data_type=cache.dtype)
    o, lse = w.run(q.to(DEV), cache.to(DEV), return_lse=True, q_len_per_req=n)
    assert o.shape == q.shape and lse.shape == (len(lens) * n, hq)
    torch.cuda.synchronize()
    k4, v4 = orc.unpack_cache(cache, "NHD")
    rope = (1.0, 1e4) if pos_mode == "ROPE_LLAMA" else None
    for b, L in enumerate(lens):
        k, v = orc.gather_paged_kv(k4, v4, indptr, indices, last, b, "NHD")
        for t in range(n):
            Lt = max(L - (n - 1 - t), 0)
            o_ref, lse_ref = orc.decode_attention_ref(q[b * n + t], k[:Lt].float(), v[:Lt].float(), rope=rope)
            assert (o[b * n + t].float().cpu() - o_ref).abs().max().item() <= (O_TOL_FP8 if is8 else O_TOL)
            assert (lse[b * n + t].cpu() - lse_ref).abs().max().item() <= LSE_TOL
    # second call reuses the plan of the expanded table; a wrong row count is refused
    o2 = w.run(q.to(DEV), cache.to(DEV), q_len_per_req=n)
    assert torch.equal(o2, o)
    with pytest.raises(ValueError):
        w.run(q[:-1].to(DEV), cache.to(DEV), q_len_per_req=n)


@pytest.mark.parametrize("n", [2, 3, 4, 8])
@pytest.mark.parametrize("hq,hkv", [(8, 8), (8, 4), (16, 4), (32, 4), (24, 2)])
@pytest.mark.parametrize("d,page,layout", [(128, 16, "NHD"), (128, 1, "NHD"), (64, 16, "HND"), (256, 8, "NHD")])
@pytest.mark.parametrize("variant", ["plain", "alibi+cap", "sinks"])
def test_multi_token_rows_share_the_kv_stream(n, hq, hkv, d, page, layout, variant):
    """run(q_len_per_req = n) with 16-bit KV: the n rows of a request are extra columns of its GQA group
    (FiDecodeRunArgs.q_rows; plan with n x the query heads on the request's own table) — every (row, head)
    against the oracle on the row's causal prefix, for groups 1 .. 12 (virtual KV heads above 8 columns),
    both loaders, split units, empty rows (kv_len < n), ALiBi + soft cap, sinks; the per-row-request path
    (FI_B200_MTP_ROWS=0) gives the same."""
    import os

    from flashinfer_b200.utils import get_alibi_slopes

    fi = _fi()
    lens = [5, 97, 3000, 2, 64, 1, 7000]
    cache, indptr, indices, last = make_paged_case(lens, hkv, d, page, layout, torch.float32, seed=141)
    cache = cache.to(torch.float16)
    q = make_q(len(lens) * n, hq, d, torch.float16, seed=142)
    kw, okw, rkw = {}, {}, {}
    if variant == "alibi+cap":
        kw = dict(pos_encoding_mode="ALIBI", logits_soft_cap=30.0)
        okw = dict(alibi_slopes=get_alibi_slopes(hq), logits_soft_cap=30.0)
    elif variant == "sinks":
        rkw = dict(sinks=torch.linspace(-1.0, 2.0, hq, device=DEV))
        okw = dict(sinks=rkw["sinks"].cpu())
    w = fi.BatchDecodeWithPagedKVCacheWrapper(_workspace(), layout)
    w.plan(indptr, indices, last, hq, hkv, d, page, q_data_type=torch.float16, kv_data_type=cache.dtype, **kw)
    o, lse = w.run(q.to(DEV), cache.to(DEV), return_lse=True, q_len_per_req=n, **rkw)
    assert w._mtp_children[n]._q_rows == n and o.shape == q.shape and lse.shape == (len(lens) * n, hq)
    torch.cuda.synchronize()
    k4, v4 = orc.unpack_cache(cache, layout)
    for b, L in enumerate(lens):
        k, v = orc.gather_paged_kv(k4, v4, indptr, indices, last, b, layout)
        for t in range(n):
            Lt = max(L - (n - 1 - t), 0)
            o_ref, lse_ref = orc.decode_attention_ref(q[b * n + t], k[:Lt].float(), v[:Lt].float(), **okw)
            assert (o[b * n + t].float().cpu() - o_ref).abs().max().item() <= O_TOL, (b, t)
            if variant == "alibi+cap":
                assert _alibi_lse_ok(lse[b * n + t].cpu(), lse_ref)
            elif Lt > 0 or variant != "sinks":
                assert (lse[b * n + t].cpu() - lse_ref).abs().max().item() <= LSE_TOL, (b, t)
    if variant == "plain" and n in (2, 4):
        os.environ["FI_B200_MTP_ROWS"] = "0"
        try:
            w2 = fi.BatchDecodeWithPagedKVCacheWrapper(_workspace(), layout)
            w2.plan(indptr, indices, last, hq, hkv, d, page, q_data_type=torch.float16, kv_data_type=cache.dtype)
            o2 = w2.run(q.to(DEV), cache.to(DEV), q_len_per_req=n)
            assert w2._mtp_children[n]._q_rows == 1
        finally:
            del os.environ["FI_B200_MTP_ROWS"]
        assert (o2.float() - o.float()).abs().max().item() <= 4e-3


@pytest.mark.parametrize("hq,hkv", [(16, 1), (12, 1), (40, 4), (128, 8), (36, 4), (24, 1)])
@pytest.mark.parametrize("d,page,layout", [(128, 16, "NHD"), (64, 32, "HND"), (128, 8, "NHD")])
@pytest.mark.parametrize("variant", ["plain", "window+cap", "alibi", "sinks"])
def test_gqa_groups_of_9_to_16_are_one_unit(hq, hkv, d, page, layout, variant):
    """GQA groups of 9 .. 16 (Llama-3-405B: 128 / 8) with 16-bit KV: one unit of up to 16 columns
    (FI_PLAN_WIDE_UNITS, two MMA N tiles per KV pass) instead of two virtual KV heads; group 24 splits
    into two units of 12.  Oracle parity incl. split units, empty requests, the logits variants; the
    plan without the flag (FI_B200_WIDE_UNITS=0) gives the same; ROPE_LLAMA through forward() re-plans."""
    import os

    from flashinfer_b200.utils import get_alibi_slopes

    fi = _fi()
    lens = [1, 97, 3000, 0, 64, 9000, 513]
    cache, indptr, indices, last = make_paged_case(lens, hkv, d, page, layout, torch.float32, seed=151)
    cache = cache.to(torch.float16)
    q = make_q(len(lens), hq, d, torch.float16, seed=152)
    kw, okw, rkw = {}, {}, {}
    if variant == "window+cap":
        kw, okw = dict(window_left=700, logits_soft_cap=30.0), dict(window_left=700, logits_soft_cap=30.0)
    elif variant == "alibi":
        kw, okw = dict(pos_encoding_mode="ALIBI"), dict(alibi_slopes=get_alibi_slopes(hq))
    elif variant == "sinks":
        rkw = dict(sinks=torch.linspace(-1.0, 2.0, hq, device=DEV))
    w = fi.BatchDecodeWithPagedKVCacheWrapper(_workspace(), layout)
    w.plan(indptr, indices, last, hq, hkv, d, page, q_data_type=torch.float16, kv_data_type=cache.dtype, **kw)
    group = hq // hkv
    assert int(w._plan_info[13]) == hkv * (1 if group <= 16 else 2)  # (virtual) KV heads of the plan
    o, lse = w.run(q.to(DEV), cache.to(DEV), return_lse=True, **rkw)
    torch.cuda.synchronize()
    if variant == "sinks":
        for i in (1, 2, 5):
            k, v = orc.gather_paged_kv(*orc.unpack_cache(cache, layout), indptr, indices, last, i, layout)
            o_ref, _ = orc.decode_attention_ref(q[i], k.float(), v.float(), sinks=rkw["sinks"].cpu())
            assert (o[i].float().cpu() - o_ref).abs().max().item() <= O_TOL
    else:
        o_ref, lse_ref = orc.batch_decode_ref(q, cache, indptr, indices, last, layout, **okw)
        assert (o.float().cpu() - o_ref).abs().max().item() <= O_TOL
        if variant == "alibi":
            assert _alibi_lse_ok(lse.cpu(), lse_ref)
        else:
            assert (lse.cpu() - lse_ref).abs().max().item() <= LSE_TOL
    if variant in ("plain", "window+cap"):
        os.environ["FI_B200_WIDE_UNITS"] = "0"
        try:
            w2 = fi.BatchDecodeWithPagedKVCacheWrapper(_workspace(), layout)
            w2.plan(indptr, indices, last, hq, hkv, d, page, q_data_type=torch.float16, kv_data_type=cache.dtype, **kw)
            assert int(w2._plan_info[13]) > int(w._plan_info[13])
            o2 = w2.run(q.to(DEV), cache.to(DEV))
        finally:
            del os.environ["FI_B200_WIDE_UNITS"]
        assert (o2.float() - o.float()).abs().max().item() <= 4e-3
    if variant == "plain":
        for i in (1, 2, 5):  # single decode takes the same 16-column units (request 1: shorter than a tile, it does not)
            k, v = orc.gather_paged_kv(*orc.unpack_cache(cache, layout), indptr, indices, last, i, layout)
            kd, vd = (k, v) if layout == "NHD" else (k.transpose(0, 1), v.transpose(0, 1))
            o1, l1 = fi.single_decode_with_kv_cache(q[i].to(DEV), kd.contiguous().to(DEV), vd.contiguous().to(DEV),
                                                    kv_layout=layout, return_lse=True)
            assert (o1.float().cpu() - o_ref[i]).abs().max().item() <= O_TOL
            assert (l1.cpu() - lse_ref[i]).abs().max().item() <= LSE_TOL
        o3 = w.forward(q.to(DEV), cache.to(DEV), pos_encoding_mode="ROPE_LLAMA")  # FI_ERR_REPLAN underneath
        o3_ref, _ = orc.batch_decode_ref(q, cache, indptr, indices, last, layout, rope=(1.0, 1e4))
        assert (o3.float().cpu() - o3_ref).abs().max().item() <= O_TOL


@pytest.mark.parametrize("kv_dtype", [torch.float8_e4m3fn, torch.float8_e5m2])
@pytest.mark.parametrize("n,hq,hkv", [(1, 128, 8), (1, 12, 1), (1, 40, 4), (2, 8, 8), (2, 16, 4), (2, 32, 4), (3, 16, 4),
                                      (4, 32, 8), (4, 8, 1), (8, 24, 2)])
@pytest.mark.parametrize("page,layout", [(16, "NHD"), (8, "HND"), (64, "NHD")])
@pytest.mark.parametrize("variant", ["plain", "alibi+cap", "sinks"])
def test_fp8_wide_units(kv_dtype, n, hq, hkv, page, layout, variant):
    """fp8 KV on the tcgen05 kernel's wide mode: up to 16 (row, head) columns per unit — GQA groups of
    9 .. 16 as one unit (n = 1) and run(q_len_per_req = n) with the rows sharing the KV stream — against
    the oracle on every row's causal prefix; split units, empty rows, ALiBi + soft cap, sinks."""
    from flashinfer_b200.utils import get_alibi_slopes

    fi = _fi()
    d = 128
    lens = [5, 97, 3000, 2, 64, 1, 7000, 0]
    cache, indptr, indices, last = make_paged_case(lens, hkv, d, page, layout, torch.float32, seed=161)
    cache = (cache * 0.5).to(kv_dtype)
    q = make_q(len(lens) * n, hq, d, torch.float16, seed=162)
    kw, okw, rkw = {}, {}, {}
    if variant == "alibi+cap":
        kw = dict(pos_encoding_mode="ALIBI", logits_soft_cap=30.0)
        okw = dict(alibi_slopes=get_alibi_slopes(hq), logits_soft_cap=30.0)
    elif variant == "sinks":
        rkw = dict(sinks=torch.linspace(-1.0, 2.0, hq, device=DEV))
        okw = dict(sinks=rkw["sinks"].cpu())
    w = fi.BatchDecodeWithPagedKVCacheWrapper(_workspace(), layout)
    w.plan(indptr, indices, last, hq, hkv, d, page, q_data_type=torch.float16, kv_data_type=kv_dtype, **kw)
    if n == 1:
        assert int(w._plan_info[13]) == hkv  # one unit per KV head: no virtual heads
        o, lse = w.run(q.to(DEV), cache.to(DEV), return_lse=True, k_scale=0.5, v_scale=2.0, **rkw)
    else:
        o, lse = w.run(q.to(DEV), cache.to(DEV), return_lse=True, q_len_per_req=n, k_scale=0.5, v_scale=2.0, **rkw)
        assert w._mtp_children[n]._q_rows == n
    torch.cuda.synchronize()
    k4, v4 = orc.unpack_cache(cache, layout)
    for b, L in enumerate(lens):
        k, v = orc.gather_paged_kv(k4, v4, indptr, indices, last, b, layout)
        for t in range(n):
            Lt = max(L - (n - 1 - t), 0)
            o_ref, lse_ref = orc.decode_attention_ref(q[b * n + t], k[:Lt].float(), v[:Lt].float(),
                                                      0.5 / math.sqrt(d), **okw)
            assert (o[b * n + t].float().cpu() - 2.0 * o_ref).abs().max().item() <= O_TOL_FP8, (b, t)
            if variant == "alibi+cap":
                assert _alibi_lse_ok(lse[b * n + t].cpu(), lse_ref)
            elif Lt > 0 or variant != "sinks":
                assert (lse[b * n + t].cpu() - lse_ref).abs().max().item() <= LSE_TOL, (b, t)


@pytest.mark.parametrize("kv_dtype", [torch.float8_e4m3fn, torch.float8_e5m2])
@pytest.mark.parametrize("layout", ["NHD", "HND"])
@pytest.mark.parametrize("pos_mode", ["NONE", "ROPE_LLAMA"])
def test_fp8_kv_at_head_dim_64(kv_dtype, layout, pos_mode):
    """fp8 KV at head_dim 64 (the reference dispatches it, include/flashinfer/utils.cuh:175-202): 64-byte rows
    run on the CUDA-core variant kernel; batch and single decode against the oracle, split units included."""
    fi = _fi()
    hq, hkv, d, page = 16, 4, 64, 16
    lens = [1, 54, 97, 512, 9000, 0, 4099, 777]
    cache, indptr, indices, last = make_paged_case(lens, hkv, d, page, layout, torch.float32, seed=91)
    cache = (cache * 0.5).to(kv_dtype)
    q = make_q(len(lens), hq, d, torch.bfloat16, seed=92)
    k_scale, v_scale = 0.5, 0.25
    w = fi.BatchDecodeWithPagedKVCacheWrapper(_workspace(), layout)
    w.plan(indptr, indices, last, hq, hkv, d, page, pos_encoding_mode=pos_mode, q_data_type=torch.bfloat16,
           kv_data_type=kv_dtype)
    o, lse = w.run(q.to(DEV), cache.to(DEV), k_scale=k_scale, v_scale=v_scale, return_lse=True)
    rope = (1.0, 1e4) if pos_mode == "ROPE_LLAMA" else None
    o_ref, lse_ref = orc.batch_decode_ref(q, cache, indptr, indices, last, layout, k_scale=k_scale,
                                          v_scale=v_scale, rope=rope)
    assert (o.float().cpu() - o_ref).abs().max().item() <= O_TOL_FP8
    assert (lse.cpu() - lse_ref).abs().max().item() <= LSE_TOL
    # single decode on request 4 (9000 tokens: the implicit split + merge kernel)
    k4, v4 = orc.unpack_cache(cache, layout)
    k, v = orc.gather_paged_kv(k4, v4, indptr, indices, last, 4, layout)
    kd, vd = (k, v) if layout == "NHD" else (k.transpose(0, 1), v.transpose(0, 1))
    o1, l1 = fi.single_decode_with_kv_cache(q[4].to(DEV), kd.contiguous().to(DEV), vd.contiguous().to(DEV),
                                            kv_layout=layout, pos_encoding_mode=pos_mode, k_scale=k_scale,
                                            v_scale=v_scale, return_lse=True)
    assert (o1.float().cpu() - o_ref[4]).abs().max().item() <= O_TOL_FP8
    assert (l1.cpu() - lse_ref[4]).abs().max().item() <= LSE_TOL
    if pos_mode == "NONE":
        w.plan(indptr, indices, last, hq, hkv, d, page, pos_encoding_mode="ALIBI", q_data_type=torch.bfloat16,
               kv_data_type=kv_dtype)
        with pytest.raises(RuntimeError):
            w.run(q.to(DEV), cache.to(DEV))


@pytest.mark.parametrize("kv_dtype", [torch.float8_e4m3fn, torch.float8_e5m2])
@pytest.mark.parametrize("page,hq,hkv", [(16, 64, 8), (8, 16, 2), (32, 8, 8), (64, 12, 4), (128, 32, 4)])
@pytest.mark.parametrize("variant", ["plain", "window+cap", "alibi", "sinks"])
def test_fp8_head_pair_plans(kv_dtype, page, hq, hkv, variant):
    """fp8 KV in an NHD cache: the wrapper plans (request, KV-head pair) units for the tcgen05 kernel
    (FI_PLAN_HEAD_PAIRS, include/fi_b200.h; plan-info slot 31).  Oracle parity across page sizes, GQA
    groups 1 / 3 / 8, split units, empty and one-token requests and the logits variants."""
    fi = _fi()
    from flashinfer_b200.utils import get_alibi_slopes

    d = 128
    lens = [1, 54, 97, 512, 9000, 0, 4099, 777, 64, 65, 20000]
    cache, indptr, indices, last = make_paged_case(lens, hkv, d, page, "NHD", torch.float32, seed=95)
    cache = (cache * 0.5).to(kv_dtype)
    q = make_q(len(lens), hq, d, torch.bfloat16, seed=96)
    kw, okw, rkw = {}, {}, {}
    if variant == "window+cap":
        kw, okw = dict(window_left=700, logits_soft_cap=30.0), dict(window_left=700, logits_soft_cap=30.0)
    elif variant == "alibi":
        kw, okw = dict(pos_encoding_mode="ALIBI"), dict(alibi_slopes=get_alibi_slopes(hq))
    w = fi.BatchDecodeWithPagedKVCacheWrapper(_workspace(), "NHD")
    w.plan(indptr, indices, last, hq, hkv, d, page, q_data_type=torch.bfloat16, kv_data_type=kv_dtype, **kw)
    assert int(w._plan_info[31]) == 1 and int(w._plan_info[13]) == hkv // 2
    assert int(w._plan_info[3]) > 0  # split units: the fused merge runs on pair partials
    sinks = None
    if variant == "sinks":
        sinks = torch.linspace(-2.0, 3.0, hq, device=DEV)
    o, lse = w.run(q.to(DEV), cache.to(DEV), k_scale=0.5, v_scale=0.25, return_lse=True, sinks=sinks)
    torch.cuda.synchronize()
    if variant == "sinks":
        for i in (1, 4, 7, 10):
            k, v = orc.gather_paged_kv(*orc.unpack_cache(cache, "NHD"), indptr, indices, last, i, "NHD")
            o_ref, _ = orc.decode_attention_ref(q[i], k.float(), v.float(), 0.5 / math.sqrt(d), sinks=sinks.cpu())
            assert (o[i].float().cpu() - o_ref * 0.25).abs().max().item() <= O_TOL_FP8
        return
    o_ref, lse_ref = orc.batch_decode_ref(q, cache, indptr, indices, last, "NHD", k_scale=0.5, v_scale=0.25, **okw)
    assert (o.float().cpu() - o_ref).abs().max().item() <= O_TOL_FP8
    if variant == "alibi":
        assert _alibi_lse_ok(lse.cpu(), lse_ref)
    else:
        assert (lse.cpu() - lse_ref).abs().max().item() <= LSE_TOL


@pytest.mark.parametrize("q_dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("kv_dtype", [None, torch.float8_e4m3fn, torch.float8_e5m2])
@pytest.mark.parametrize("hq,hkv", [(4, 4), (8, 2), (16, 2)])
@pytest.mark.parametrize("variant", ["plain", "rope", "window+cap"])
def test_head_dim_512(q_dtype, kv_dtype, hq, hkv, variant):
    """head_dim 512 (the reference dispatches it, include/flashinfer/utils.cuh:192): runs on the CUDA-core
    variant kernel, groups above 4 as virtual KV heads; batch (split units, merge) and single decode
    against the oracle; ALiBi is refused."""
    fi = _fi()
    d, page, layout = 512, 16, "NHD"
    lens = [1, 130, 3000, 0, 64, 9000]
    is8 = kv_dtype is not None
    cache, indptr, indices, last = make_paged_case(lens, hkv, d, page, layout, torch.float32, seed=111)
    cache = (cache * 0.5 if is8 else cache).to(kv_dtype or q_dtype)
    q = make_q(len(lens), hq, d, q_dtype, seed=112)
    kw, okw = {}, {}
    if variant == "rope":
        kw, okw = dict(pos_encoding_mode="ROPE_LLAMA"), dict(rope=(1.0, 1e4))
    elif variant == "window+cap":
        kw, okw = dict(window_left=500, logits_soft_cap=20.0), dict(window_left=500, logits_soft_cap=20.0)
    w = fi.BatchDecodeWithPagedKVCacheWrapper(_workspace(), layout)
    w.plan(indptr, indices, last, hq, hkv, d, page, q_data_type=q_dtype, kv_data_type=cache.dtype, **kw)
    o, lse = w.run(q.to(DEV), cache.to(DEV), return_lse=True)
    o_ref, lse_ref = orc.batch_decode_ref(q, cache, indptr, indices, last, layout, **okw)
    assert (o.float().cpu() - o_ref).abs().max().item() <= (O_TOL_FP8 if is8 else O_TOL)
    assert (lse.cpu() - lse_ref).abs().max().item() <= LSE_TOL
    if variant != "window+cap":
        k, v = orc.gather_paged_kv(*orc.unpack_cache(cache, layout), indptr, indices, last, 2, layout)
        o1 = fi.single_decode_with_kv_cache(q[2].to(DEV), k.contiguous().to(DEV), v.contiguous().to(DEV),
                                            pos_encoding_mode=kw.get("pos_encoding_mode", "NONE"))
        assert (o1.float().cpu() - o_ref[2]).abs().max().item() <= (O_TOL_FP8 if is8 else O_TOL)
    if variant == "plain":
        w.plan(indptr, indices, last, hq, hkv, d, page, q_data_type=q_dtype, kv_data_type=cache.dtype,
               pos_encoding_mode="ALIBI")
        with pytest.raises(RuntimeError):
            w.run(q.to(DEV), cache.to(DEV))


@pytest.mark.parametrize("kv_dtype", [None, torch.float8_e4m3fn])
def test_run_into_a_pinned_host_tensor(kv_dtype):
    """run(out = pinned host tensor): the epilogue (direct and merged units) stores o over PCIe; bit-equal
    to the device result after a stream sync.  Pageable host memory is refused."""
    fi = _fi()
    hq, hkv, d, page = 32, 8, 128, 16
    lens = [1, 500, 9000, 0, 4099, 64, 20000]
    cache, indptr, indices, last = make_paged_case(lens, hkv, d, page, "NHD", torch.float32, seed=101)
    cache = (cache * 0.5 if kv_dtype else cache).to(kv_dtype or torch.bfloat16)
    q = make_q(len(lens), hq, d, torch.bfloat16, seed=102).to(DEV)
    w = fi.BatchDecodeWithPagedKVCacheWrapper(_workspace(), "NHD")
    w.plan(indptr, indices, last, hq, hkv, d, page, q_data_type=torch.bfloat16, kv_data_type=cache.dtype)
    assert int(w._plan_info[3]) > 0
    cache_d = cache.to(DEV)
    o_dev = w.run(q, cache_d)
    o_host = torch.full(q.shape, float("nan"), dtype=q.dtype).pin_memory()
    ret = w.run(q, cache_d, out=o_host)
    assert ret is o_host
    torch.cuda.synchronize()
    assert torch.equal(o_host, o_dev.cpu())
    with pytest.raises(ValueError):
        w.run(q, cache_d, out=torch.empty(q.shape, dtype=q.dtype))


@pytest.mark.parametrize("kv_dtype", [None, torch.float8_e4m3fn])
def test_step_is_plan_plus_run(kv_dtype):
    """wrapper.step(table, q, cache): plan + run through ONE library call (fi_batch_decode_step) with the
    problem description of the last plan() and the scales of the last run() — bit-equal to plan() + run()
    for new page tables and batch sizes, into device or pinned host memory; device tables and a
    different cache tensor take the two-call path."""
    fi = _fi()
    hq, hkv, d, page = 32, 8, 128, 16
    is8 = kv_dtype is not None
    lens_all = [[300, 5000, 17, 0, 9000], [301, 5001, 18, 1, 9001], [64, 128, 7000], [12000, 1, 2, 3, 4, 5, 6]]
    big = max(sum(-(-l // page) for l in lens) for lens in lens_all) + 8
    g = torch.Generator().manual_seed(121)
    cache = torch.randn(big, 2, page, hkv, d, generator=g)
    cache = (cache * 0.5 if is8 else cache).to(kv_dtype or torch.bfloat16).to(DEV)
    w = fi.BatchDecodeWithPagedKVCacheWrapper(_workspace(), "NHD")
    ref = fi.BatchDecodeWithPagedKVCacheWrapper(torch.empty(64 << 20, dtype=torch.uint8, device=DEV), "NHD")
    kw = dict(q_data_type=torch.bfloat16, kv_data_type=cache.dtype, logits_soft_cap=30.0)
    rkw = dict(k_scale=0.5, v_scale=2.0)

    def table(lens, seed):
        pages = [-(-l // page) for l in lens]
        perm = torch.randperm(big, generator=torch.Generator().manual_seed(seed))[: sum(pages)].to(torch.int32)
        indptr = torch.tensor([0] + list(np.cumsum(pages)), dtype=torch.int32)
        last = torch.tensor([(l - 1) % page + 1 if l else 0 for l in lens], dtype=torch.int32)
        return indptr, perm.contiguous(), last

    with pytest.raises(RuntimeError):
        w.step(*table(lens_all[0], 1), make_q(5, hq, d, torch.bfloat16).to(DEV), cache)  # no plan() yet
    for it, lens in enumerate(lens_all):
        indptr, indices, last = table(lens, 10 + it)
        q = make_q(len(lens), hq, d, torch.bfloat16, seed=130 + it).to(DEV)
        ref.plan(indptr, indices, last, hq, hkv, d, page, **kw)
        o_ref = ref.run(q, cache, **rkw)
        if it == 0:  # what a serving loop has done before its first step()
            w.plan(indptr, indices, last, hq, hkv, d, page, **kw)
            w.run(q, cache, **rkw)
        if it == 2:
            o = torch.empty(q.shape, dtype=q.dtype).pin_memory()
            assert w.step(indptr, indices, last, q, cache, out=o) is o
        elif it == 3:  # device tables: plan() + run() underneath
            o = w.step(indptr.to(DEV), indices.to(DEV), last.to(DEV), q, cache)
        else:
            o = w.step(indptr, indices, last, q, cache)
        torch.cuda.synchronize()
        assert torch.equal(o.cpu(), o_ref.cpu()), it
        # the wrapper is left as after plan(): run() alone gives the same again
        assert torch.equal(w.run(q, cache, **rkw).cpu(), o_ref.cpu())
    other = cache.clone()
    assert torch.equal(w.step(indptr, indices, last, q, other).cpu(), o_ref.cpu())


def test_fp8_head_pair_plan_falls_back_when_run_needs_rope():
    """A head-pair plan cannot feed the RoPE variant kernel: forward(pos_encoding_mode="ROPE_LLAMA") after a
    NONE plan gets FI_ERR_REPLAN from the C-ABI, and the wrapper plans once more without the flag."""
    fi = _fi()
    hq, hkv, d, page = 16, 4, 128, 16
    lens = [300, 5000, 17]
    cache, indptr, indices, last = make_paged_case(lens, hkv, d, page, "NHD", torch.float32, seed=97)
    cache = (cache * 0.5).to(torch.float8_e4m3fn)
    q = make_q(len(lens), hq, d, torch.float16, seed=98)
    w = fi.BatchDecodeWithPagedKVCacheWrapper(_workspace(), "NHD")
    w.plan(indptr, indices, last, hq, hkv, d, page, q_data_type=torch.float16, kv_data_type=cache.dtype)
    assert int(w._plan_info[31]) == 1
    o = w.forward(q.to(DEV), cache.to(DEV), pos_encoding_mode="ROPE_LLAMA")
    assert int(w._plan_info[31]) == 0 and not w._head_pairs_ok
    o_ref, _ = orc.batch_decode_ref(q, cache, indptr, indices, last, "NHD", rope=(1.0, 1e4))
    assert (o.float().cpu() - o_ref).abs().max().item() <= O_TOL_FP8


@pytest.mark.parametrize("kv_dtype", [None, torch.float8_e4m3fn])
def test_multi_token_decode_in_a_cuda_graph_with_replans(kv_dtype):
    """run(q_len_per_req = n) captured in a CUDA graph (the reference captures its tensor-core decode,
    tests/attention/test_tensor_cores_decode.py:200-330): warm-up call outside the capture, then
    replays after re-plans with different kv lengths — the per-row child plan follows every plan()."""
    fi = _fi()
    hq, hkv, d, page, n = 16, 4, 128, 16, 3
    lens_a, lens_b = [100, 2000, 31, 977, 5], [4000, 7, 64, 1500, 333]
    is8 = kv_dtype is not None
    cache, ip_b, ix_b, lp_b = make_paged_case(lens_b, hkv, d, page, "NHD", torch.float32, seed=81)
    cache = (cache * 0.5 if is8 else cache).to(kv_dtype or torch.float16)
    _, ip_a, ix_a, lp_a = make_paged_case(lens_a, hkv, d, page, "NHD", torch.float16, seed=82)
    bs = len(lens_a)
    max_pages = 4096
    w = fi.CUDAGraphBatchDecodeWithPagedKVCacheWrapper(
        _workspace(), torch.zeros(bs + 1, dtype=torch.int32, device=DEV),
        torch.zeros(max_pages, dtype=torch.int32, device=DEV), torch.zeros(bs, dtype=torch.int32, device=DEV), "NHD")
    q = make_q(bs * n, hq, d, torch.float16, seed=83).to(DEV)
    cache_d = cache.to(DEV)
    out = torch.empty_like(q)
    kw = dict(q_data_type=torch.float16, kv_data_type=cache.dtype)
    w.plan(ip_a, ix_a, lp_a, hq, hkv, d, page, **kw)
    w.run(q, cache_d, out=out, q_len_per_req=n)  # warm-up: creates and plans the per-row child
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        w.run(q, cache_d, out=out, q_len_per_req=n)
    k4, v4 = orc.unpack_cache(cache, "NHD")
    for lens, ip, ix, lp in ((lens_a, ip_a, ix_a, lp_a), (lens_b, ip_b, ix_b, lp_b), (lens_a, ip_a, ix_a, lp_a)):
        w.plan(ip, ix, lp, hq, hkv, d, page, **kw)
        graph.replay()
        torch.cuda.synchronize()
        for b, L in enumerate(lens):
            k, v = orc.gather_paged_kv(k4, v4, ip, ix, lp, b, "NHD")
            for t in range(n):
                Lt = max(L - (n - 1 - t), 0)
                o_ref, _ = orc.decode_attention_ref(q[b * n + t].cpu(), k[:Lt].float(), v[:Lt].float())
                assert (out[b * n + t].float().cpu() - o_ref).abs().max().item() <= (O_TOL_FP8 if is8 else O_TOL)


@pytest.mark.parametrize("kv_dtype,pos_mode,hq", [(None, "NONE", 16), (torch.float8_e4m3fn, "NONE", 32),
                                                  (None, "ROPE_LLAMA", 16), (torch.float8_e4m3fn, "ROPE_LLAMA", 16),
                                                  (None, "NONE", 64)])
def test_attention_sinks(kv_dtype, pos_mode, hq):
    """run(sinks=[Hq]) (reference decode.py:1176; oracle = the reference's sink_softmax,
    tests/test_helpers/sink_attention_reference.py:23-36): one extra logit per query head in the
    softmax denominator; split requests add it once, in the merge."""
    fi = _fi()
    hkv, d, page = 4, 128, 16
    lens = [1, 97, 512, 9000, 0, 64, 2048]
    is8 = kv_dtype is not None
    cache, indptr, indices, last = make_paged_case(lens, hkv, d, page, "NHD", torch.float32, seed=81)
    cache = (cache * 0.5 if is8 else cache).to(kv_dtype or torch.float16)
    q = make_q(len(lens), hq, d, torch.float16, seed=82)
    g = torch.Generator().manual_seed(83)
    sinks = torch.randn(hq, generator=g) * 2 + 1
    w = fi.BatchDecodeWithPagedKVCacheWrapper(_workspace(), "NHD")
    w.plan(indptr, indices, last, hq, hkv, d, page, pos_encoding_mode=pos_mode, q_data_type=torch.float16,
           kv_data_type=cache.dtype)
    assert int(w._plan_info[3]) > 0  # some unit is split: its sink is added in the merge
    o, lse = w.run(q.to(DEV), cache.to(DEV), return_lse=True, sinks=sinks.to(DEV))
    torch.cuda.synchronize()
    k4, v4 = orc.unpack_cache(cache, "NHD")
    rope = (1.0, 1e4) if pos_mode == "ROPE_LLAMA" else None
    for b in range(len(lens)):
        k, v = orc.gather_paged_kv(k4, v4, indptr, indices, last, b, "NHD")
        o_ref, lse_ref = orc.decode_attention_ref(q[b], k.float(), v.float(), rope=rope, sinks=sinks)
        assert (o[b].float().cpu() - o_ref).abs().max().item() <= (O_TOL_FP8 if is8 else O_TOL), b
        if lens[b] > 0:
            assert (lse[b].cpu() - lse_ref).abs().max().item() <= LSE_TOL, b
    # without sinks the same wrapper gives the plain result again (the pointer is per call)
    o0 = w.run(q.to(DEV), cache.to(DEV))
    o_ref0, _ = orc.batch_decode_ref(q, cache, indptr, indices, last, "NHD", rope=rope)
    assert (o0.float().cpu() - o_ref0).abs().max().item() <= (O_TOL_FP8 if is8 else O_TOL)


def _sink_golden():
    from pathlib import Path

    z = np.load(Path(__file__).resolve().parent / "golden" / "sink_golden.npz")
    return z, [str(n) for n in z["names"]]


@pytest.mark.parametrize("name", _sink_golden()[1])
def test_attention_sinks_against_reference_golden(name):
    """batch decode with sinks (and sliding window) against outputs of the reference's own
    sink_attention_unified (tests/golden/make_golden_sink.py)."""
    fi = _fi()
    z, _ = _sink_golden()
    B, L, Hq, Hkv, D, win, is_bf16 = [int(x) for x in z[f"{name}/meta"]]
    dt = torch.bfloat16 if is_bf16 else torch.float16
    q, k, v, o_gold = (torch.from_numpy(z[f"{name}/{key}"].copy()).view(dt) for key in ("q", "k", "v", "o"))
    q, o_gold = q.view(B, Hq, D), o_gold.view(B, Hq, D)
    k, v = k.view(B, L, Hkv, D), v.view(B, L, Hkv, D)
    sink = torch.from_numpy(z[f"{name}/sink"].copy())
    page = 16
    npg = -(-L // page)
    cache = torch.zeros(B * npg, 2, page, Hkv, D, dtype=dt)
    for b in range(B):
        pad = torch.zeros(npg * page, Hkv, D, dtype=dt)
        pad[:L] = k[b]
        cache[b * npg:(b + 1) * npg, 0] = pad.view(npg, page, Hkv, D)
        pad = torch.zeros(npg * page, Hkv, D, dtype=dt)
        pad[:L] = v[b]
        cache[b * npg:(b + 1) * npg, 1] = pad.view(npg, page, Hkv, D)
    indptr = torch.arange(B + 1, dtype=torch.int32) * npg
    indices = torch.arange(B * npg, dtype=torch.int32)
    last = torch.full((B,), (L - 1) % page + 1, dtype=torch.int32)
    w = fi.BatchDecodeWithPagedKVCacheWrapper(_workspace(), "NHD")
    w.plan(indptr, indices, last, Hq, Hkv, D, page, q_data_type=dt, window_left=win)
    o = w.run(q.to(DEV), cache.to(DEV), sinks=sink.to(DEV))
    assert (o.float().cpu() - o_gold.float()).abs().max().item() <= O_TOL


@pytest.mark.parametrize("switches", [
    {"FI_B200_FUSED_MERGE": "0"},                              # split-KV merge as a second launch
    {"FI_B200_FP8_TC5": "0", "FI_B200_L2_HINTS": "0"},         # fp8 on the register path, no L2 hints
    {"FI_B200_HEAD_GROUP": "0", "FI_B200_PDL_LATE": "1"},      # request-major order, late PDL trigger
    {"FI_B200_ROPE_CUDA_CORE": "1"},                           # RoPE on the CUDA-core variant kernel
])
def test_ab_switches_keep_parity(switches):
    """The A/B switches of DESIGN.md §6b select older / alternative code paths that stay in the library;
    most are read once per process, so the parity tests run again in a child process with each set."""
    import os
    import subprocess
    import sys

    if os.environ.get("FI_B200_TEST_CHILD") == "1":
        pytest.skip("already inside the child run")
    env = dict(os.environ, FI_B200_TEST_CHILD="1", **switches)
    sel = ("test_batch_decode_minimum_slice or test_empty_request_regression or test_batch_decode_fp8_kv or "
           "test_split_kv_long_single_request_and_determinism or test_cuda_graph_replay_with_replan or "
           "test_batch_decode_rope or test_attention_sinks or test_batch_decode_large_gqa_groups or "
           "test_window_left_and_soft_cap")
    r = subprocess.run([sys.executable, "-m", "pytest", __file__, "-x", "-q", "-m", "gpu", "-k", sel],
                       env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert " passed" in r.stdout
