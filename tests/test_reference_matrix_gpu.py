"""GPU: the reference's own batch-decode test matrix, written the way the reference writes it
(tests/attention/test_batch_decode_kernels.py:58-205 `test_batch_decode_with_paged_kv_cache`,
:208-365 `..._with_fast_plan`, :367-500 `..._with_tuple_paged_kv_cache`): same parameter grid —
batch sizes 12 / 17 / 128, kv_len 54 .. 16384, page sizes 1 / 8 / 16, 4 KV heads x {4, 32} query
heads, head_dim 128 / 256, NHD / HND, NONE / ROPE_LLAMA, f16 and e4m3 KV, contiguous and strided
caches (the `[:, 1, :, 1, :, 1, :, 1, :]` view of a 16x larger allocation) — and the same check:
every request of the batch equals `single_decode_with_kv_cache` on its gathered K / V
(rtol = atol = 1e-3 as the reference for 16-bit KV; fp8 KV runs the tcgen05 kernel in the batch and
may split differently in the single call, so it is held to 1e-2), plus the user-allocated
output buffer.  Every sampled request is also anchored on the oracle's fp32 attention (independent of
the kernels), o within 2e-3 + 2e-3 |o| for f16 KV and 1e-2 + 1e-2 |o| for e4m3 KV, lse within 1e-3.
"""
import pytest
import torch

from oracle import attention_ref as orc

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _fi():
    import flashinfer_b200 as fi

    return fi


@pytest.fixture(autouse=True)
def _release_big_blocks():
    """The grid mixes 100 KB and 60 GB caches: hand cached blocks back between cases so that one
    case's allocation pattern (or a failed case's traceback) cannot starve the next."""
    yield
    if torch.cuda.memory_reserved() > (8 << 30):
        torch.cuda.empty_cache()


def _make_cache(batch_size, kv_len, page_size, num_kv_heads, head_dim, kv_layout, kv_dtype, contiguous_kv):
    num_pages_per_seq = (kv_len + page_size - 1) // page_size
    total_num_pages = num_pages_per_seq * batch_size
    if kv_layout == "HND":
        kv_shape = [total_num_pages, 2, num_kv_heads, page_size, head_dim]
    else:
        kv_shape = [total_num_pages, 2, page_size, num_kv_heads, head_dim]
    if not contiguous_kv:
        tmp = [kv_shape[0]]
        for v in kv_shape[1:]:
            tmp.append(2)
            tmp.append(v)
        kv_data_fp32 = torch.randn(*tmp, dtype=torch.float32, device=DEV)
        kv_data = kv_data_fp32.to(kv_dtype)
        kv_data = kv_data[:, 1, :, 1, :, 1, :, 1, :]
        kv_data_fp32 = kv_data_fp32[:, 1, :, 1, :, 1, :, 1, :]
        assert kv_data.stride(-4) != kv_data.shape[-3] * kv_data.shape[-2] * kv_data.shape[-1]
    else:
        kv_data_fp32 = torch.randn(*kv_shape, dtype=torch.float32, device=DEV)
        kv_data = kv_data_fp32.to(kv_dtype)
    kv_indptr = torch.arange(0, batch_size + 1, device=DEV, dtype=torch.int32) * num_pages_per_seq
    kv_indices = torch.arange(0, total_num_pages, device=DEV, dtype=torch.int32)
    kv_last_page_len = torch.full((batch_size,), (kv_len - 1) % page_size + 1, dtype=torch.int32, device=DEV)
    return kv_data, kv_data_fp32, kv_indptr, kv_indices, kv_last_page_len


def _request_kv(kv_data_fp32, kv_indptr, kv_last_page_len, i, kv_layout, num_kv_heads, head_dim, kv_dtype):
    """K / V of request i exactly as the reference test gathers them (:142-175)."""
    perm_dims = [0, 2, 1, 3] if kv_layout == "HND" else [0, 1, 2, 3]
    perm_dims_last = [1, 0, 2] if kv_layout == "HND" else [0, 1, 2]
    out = []
    for kv in (0, 1):
        full = kv_data_fp32[kv_indptr[i]: kv_indptr[i + 1] - 1, kv].permute(*perm_dims)
        lastp = (kv_data_fp32[kv_indptr[i + 1] - 1, kv, :, : kv_last_page_len[i]] if kv_layout == "HND"
                 else kv_data_fp32[kv_indptr[i + 1] - 1, kv, : kv_last_page_len[i], :])
        out.append(torch.cat([full.reshape(-1, num_kv_heads, head_dim),
                              lastp.permute(*perm_dims_last).reshape(-1, num_kv_heads, head_dim)], dim=0)
                   .to(kv_dtype))
    return out


@pytest.mark.parametrize("batch_size", [12, 17, 128])
@pytest.mark.parametrize("kv_len", [54, 97, 512, 2048, 16384])
@pytest.mark.parametrize("page_size", [1, 8, 16])
@pytest.mark.parametrize("num_kv_heads", [4])
@pytest.mark.parametrize("num_qo_heads", [4, 32])
@pytest.mark.parametrize("head_dim", [128, 256])
@pytest.mark.parametrize("kv_layout", ["NHD", "HND"])
@pytest.mark.parametrize("pos_encoding_mode", ["NONE", "ROPE_LLAMA"])
@pytest.mark.parametrize("kv_dtype", [torch.float16, torch.float8_e4m3fn])
@pytest.mark.parametrize("contiguous_kv", [True, False])
def test_batch_decode_with_paged_kv_cache(batch_size, kv_len, page_size, num_kv_heads, num_qo_heads,
                                          head_dim, kv_layout, pos_encoding_mode, kv_dtype, contiguous_kv):
    fi = _fi()
    q_dtype = torch.float16
    # the whole grid (5760 cases) runs except where the reference's own construction — an fp32
    # master copy, 16x larger still for the strided variant — does not fit next to its kv_dtype copy
    # in the 180 GB of one B200 (about 4 % of the grid: strided caches of 128 x 16384 tokens, of
    # 12 / 17 x 16384 tokens at head_dim 256 and of 128 x 2048 tokens at head_dim 256)
    elems = batch_size * kv_len * num_kv_heads * head_dim * 2 * (1 if contiguous_kv else 16)
    if elems * 4 > (40 << 30):
        pytest.skip("the fp32 master copy of this strided case does not fit the GPU's memory")
    torch.manual_seed(42)
    q = torch.randn(batch_size, num_qo_heads, head_dim, device=DEV, dtype=q_dtype)
    kv_data, kv_data_fp32, kv_indptr, kv_indices, kv_last_page_len = _make_cache(
        batch_size, kv_len, page_size, num_kv_heads, head_dim, kv_layout, kv_dtype, contiguous_kv)
    workspace_buffer = torch.zeros(32 * 1024 * 1024, dtype=torch.int8, device=DEV)
    wrapper = fi.BatchDecodeWithPagedKVCacheWrapper(workspace_buffer, kv_layout)
    wrapper.plan(kv_indptr, kv_indices, kv_last_page_len, num_qo_heads, num_kv_heads, head_dim, page_size,
                 logits_soft_cap=0.0, pos_encoding_mode=pos_encoding_mode, data_type=kv_dtype,
                 q_data_type=q_dtype)
    o, lse = wrapper.run(q, kv_data, return_lse=True)
    assert lse.shape == (batch_size, num_qo_heads) and lse.dtype == torch.float32

    is8 = kv_dtype == torch.float8_e4m3fn
    tol = 1e-2 if is8 else 1e-3
    ip, lp = kv_indptr.cpu(), kv_last_page_len.cpu()
    for i in sorted({0, 1, batch_size // 2, batch_size - 1}):
        ki, vi = _request_kv(kv_data_fp32, ip, lp, i, kv_layout, num_kv_heads, head_dim, kv_dtype)
        o_ref_i = fi.single_decode_with_kv_cache(q[i], ki, vi, pos_encoding_mode=pos_encoding_mode,
                                                 logits_soft_cap=0.0)
        torch.testing.assert_close(o[i], o_ref_i, rtol=tol, atol=tol)
    # anchor EVERY sampled request on the oracle's fp32 attention (evaluated on the GPU, TF32 off;
    # request 0 also on the CPU): the batch == single check above shares kernels and math
    rope = (1.0, 1e4) if pos_encoding_mode == "ROPE_LLAMA" else None
    for i in sorted({0, 1, batch_size // 2, batch_size - 1}):
        ki, vi = _request_kv(kv_data_fp32, ip, lp, i, kv_layout, num_kv_heads, head_dim, kv_dtype)
        oi, lsei = orc.decode_attention_ref(q[i].float(), ki.float(), vi.float(), rope=rope)
        excess = (o[i].float() - oi).abs() - ((1e-2 + 1e-2 * oi.abs()) if is8 else (2e-3 + 2e-3 * oi.abs()))
        assert excess.max().item() <= 0.0, (i, excess.max().item())
        assert (lse[i] - lsei).abs().max().item() <= 1e-3
    ki, vi = _request_kv(kv_data_fp32, ip, lp, 0, kv_layout, num_kv_heads, head_dim, kv_dtype)
    o0, lse0 = orc.decode_attention_ref(q[0].cpu(), ki.float().cpu(), vi.float().cpu(), rope=rope)
    assert (o[0].float().cpu() - o0).abs().max().item() <= (3e-2 if is8 else 1e-2)
    assert (lse[0].cpu() - lse0).abs().max().item() <= 1e-3

    # test user-allocated output
    o_buffer = torch.empty_like(o)
    wrapper.run(q, kv_data, out=o_buffer)
    torch.testing.assert_close(o, o_buffer, rtol=1e-3, atol=1e-3)


@pytest.mark.parametrize("batch_size", [12, 17, 128])
@pytest.mark.parametrize("kv_len", [54, 97, 512, 2048])
@pytest.mark.parametrize("page_size", [1, 8, 16])
@pytest.mark.parametrize("num_qo_heads", [4, 32])
@pytest.mark.parametrize("head_dim", [128, 256])
@pytest.mark.parametrize("kv_dtype", [torch.float16, torch.float8_e4m3fn])
def test_batch_decode_with_paged_kv_cache_with_fast_plan(batch_size, kv_len, page_size, num_qo_heads,
                                                         head_dim, kv_dtype):
    """reference :208-365 — `fast_decode_plan` bound onto a wrapper (as SGLang does) after a
    regular warm-up plan, with device index tensors and the host copy of indptr handed in; the
    result must equal the regular plan's."""
    fi = _fi()
    from flashinfer_b200.decode import fast_decode_plan
    import functools

    num_kv_heads, kv_layout, q_dtype = 4, "NHD", torch.float16
    torch.manual_seed(7)
    q = torch.randn(batch_size, num_qo_heads, head_dim, device=DEV, dtype=q_dtype)
    kv_data, _, kv_indptr, kv_indices, kv_last_page_len = _make_cache(
        batch_size, kv_len, page_size, num_kv_heads, head_dim, kv_layout, kv_dtype, True)
    workspace_buffer = torch.zeros(32 * 1024 * 1024, dtype=torch.int8, device=DEV)
    wrapper = fi.BatchDecodeWithPagedKVCacheWrapper(workspace_buffer, kv_layout)
    wrapper.plan(kv_indptr, kv_indices, kv_last_page_len, num_qo_heads, num_kv_heads, head_dim, page_size,
                 data_type=kv_dtype, q_data_type=q_dtype)
    o_ref, lse_ref = wrapper.run(q, kv_data, return_lse=True)

    fast = fi.BatchDecodeWithPagedKVCacheWrapper(workspace_buffer.clone(), kv_layout)
    fast.plan(kv_indptr, kv_indices, kv_last_page_len, num_qo_heads, num_kv_heads, head_dim, page_size,
              data_type=kv_dtype, q_data_type=q_dtype)  # warm-up plan (reference :255-270)
    fast.begin_forward = functools.partial(fast_decode_plan, fast)
    fast.begin_forward(kv_indptr, kv_indices, kv_last_page_len, num_qo_heads, num_kv_heads, head_dim,
                       page_size, data_type=kv_dtype, q_data_type=q_dtype, non_blocking=True,
                       global_override_indptr_cpu=kv_indptr.cpu())
    o, lse = fast.run(q, kv_data, return_lse=True)
    assert torch.equal(o, o_ref) and torch.equal(lse, lse_ref)


@pytest.mark.parametrize("batch_size", [12, 17, 128])
@pytest.mark.parametrize("kv_len", [54, 97, 512, 2048])
@pytest.mark.parametrize("page_size", [1, 8, 16])
@pytest.mark.parametrize("kv_layout", ["NHD", "HND"])
@pytest.mark.parametrize("head_dim", [128, 256])
def test_batch_decode_with_tuple_paged_kv_cache(batch_size, kv_len, page_size, kv_layout, head_dim):
    """reference :367-500 — the cache handed over as a `(k_cache, v_cache)` tuple of separately
    allocated 4-D tensors equals the 5-D form."""
    fi = _fi()
    num_kv_heads, num_qo_heads, dt = 4, 32, torch.float16
    torch.manual_seed(9)
    q = torch.randn(batch_size, num_qo_heads, head_dim, device=DEV, dtype=dt)
    kv_data, _, kv_indptr, kv_indices, kv_last_page_len = _make_cache(
        batch_size, kv_len, page_size, num_kv_heads, head_dim, kv_layout, dt, True)
    workspace_buffer = torch.zeros(32 * 1024 * 1024, dtype=torch.int8, device=DEV)
    wrapper = fi.BatchDecodeWithPagedKVCacheWrapper(workspace_buffer, kv_layout)
    wrapper.plan(kv_indptr, kv_indices, kv_last_page_len, num_qo_heads, num_kv_heads, head_dim, page_size,
                 data_type=dt, q_data_type=dt)
    o5 = wrapper.run(q, kv_data)
    kv_tuple = tuple(x.squeeze(1).contiguous() for x in torch.chunk(kv_data, 2, dim=1))
    o2, lse2 = wrapper.run(q, kv_tuple, return_lse=True)
    assert torch.equal(o2, o5)
