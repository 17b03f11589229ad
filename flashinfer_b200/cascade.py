"""Attention-state merge operators — same names and semantics as the reference's
flashinfer/cascade.py (merge_state :46-99, merge_state_in_place :112-158,
merge_states :172-215), backed by csrc/merge.cu through the C-ABI.

A state is (v, s): v the normalised partial attention output, s its log-sum-exp in
**base 2** (reference include/flashinfer/attention/state.cuh:45, cascade.cuh:55-69).
"""
from __future__ import annotations

import os
from typing import Optional, Tuple

import torch

from . import _lib
from .decode import rows_can_share_the_stream
from .utils import canonicalize_torch_dtype, current_stream_ptr, device_index, dtype_code


def _check_vs(v: torch.Tensor, s: torch.Tensor, name: str) -> None:
    if v.dim() != 3:
        raise ValueError(f"{name} must be 3-D [seq_len, num_heads, head_dim]")
    if s.shape != v.shape[:2]:
        raise ValueError(f"shape of s ({list(s.shape)}) does not match {name} ({list(v.shape)})")
    if v.dtype not in (torch.float16, torch.bfloat16, torch.float32):
        raise ValueError(f"{name} must be float16 / bfloat16 / float32, got {v.dtype}")
    if s.device != v.device:
        raise ValueError("v and s must be on the same device")


def merge_state(
    v_a: torch.Tensor, s_a: torch.Tensor, v_b: torch.Tensor, s_b: torch.Tensor
) -> Tuple[torch.Tensor, torch.Tensor]:
    r"""Merge the attention states of two KV segments (reference cascade.py:46-99).

    v_* ``[seq_len, num_heads, head_dim]``, s_* ``[seq_len, num_heads]`` (base-2 LSE).
    """
    s_a = s_a.to(torch.float32).contiguous()
    s_b = s_b.to(torch.float32).contiguous()
    v_a = v_a.contiguous()
    v_b = v_b.contiguous()
    _check_vs(v_a, s_a, "v_a")
    _check_vs(v_b, s_b, "v_b")
    if v_a.shape != v_b.shape or v_a.dtype != v_b.dtype or v_a.device != v_b.device:
        raise ValueError("v_a and v_b must have the same shape, dtype and device")
    dev = v_a.device
    v_merged = torch.empty_like(v_a)
    s_merged = torch.empty_like(s_a)
    seq_len, num_heads, head_dim = v_a.shape
    _lib.check(
        _lib.load().fi_merge_state(
            v_a.data_ptr(), s_a.data_ptr(), v_b.data_ptr(), s_b.data_ptr(),
            v_merged.data_ptr(), s_merged.data_ptr(), seq_len, num_heads, head_dim,
            dtype_code(v_a.dtype), device_index(dev), current_stream_ptr(dev),
        )
    )
    return v_merged, s_merged


def merge_state_in_place(
    v: torch.Tensor,
    s: torch.Tensor,
    v_other: torch.Tensor,
    s_other: torch.Tensor,
    mask: Optional[torch.Tensor] = None,
) -> None:
    r"""Merge (v_other, s_other) into (v, s) in place (reference cascade.py:112-158).

    ``mask`` ``[seq_len]`` (bool / uint8): rows with a zero mask are left untouched
    (reference cascade.cuh:96-100).  ``s`` must be float32 for the update to be visible,
    as in the reference (which casts, and a cast copy would be discarded).
    """
    _check_vs(v, s, "v")
    _check_vs(v_other, s_other, "v_other")
    if not v.is_contiguous() or not s.is_contiguous():
        raise ValueError("v and s must be contiguous for the in-place merge")
    if s.dtype != torch.float32:
        raise ValueError("s must be float32 for merge_state_in_place")
    if v.shape != v_other.shape or v.dtype != v_other.dtype or v.device != v_other.device:
        raise ValueError("v and v_other must have the same shape, dtype and device")
    v_other = v_other.contiguous()
    s_other = s_other.to(torch.float32).contiguous()
    dev = v.device
    mask_ptr = None
    if mask is not None:
        if mask.shape != (v.shape[0],):
            raise ValueError("mask must have shape [seq_len]")
        mask = mask.to(device=dev).contiguous()
        if mask.dtype == torch.bool:
            mask = mask.view(torch.uint8)
        elif mask.dtype != torch.uint8:
            mask = (mask != 0).view(torch.uint8)
        mask_ptr = mask.data_ptr()
    seq_len, num_heads, head_dim = v.shape
    _lib.check(
        _lib.load().fi_merge_state_in_place(
            v.data_ptr(), s.data_ptr(), v_other.data_ptr(), s_other.data_ptr(), mask_ptr,
            seq_len, num_heads, head_dim, dtype_code(v.dtype), device_index(dev),
            current_stream_ptr(dev),
        )
    )


def merge_states(v: torch.Tensor, s: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    r"""Merge n attention states per position (reference cascade.py:172-215).

    v ``[seq_len, num_states, num_heads, head_dim]``, s ``[seq_len, num_states, num_heads]``
    → (``[seq_len, num_heads, head_dim]``, ``[seq_len, num_heads]``).
    """
    if v.dim() != 4:
        raise ValueError("v must be 4-D [seq_len, num_states, num_heads, head_dim]")
    if s.shape != v.shape[:3]:
        raise ValueError("shape of s does not match v")
    if v.dtype not in (torch.float16, torch.bfloat16, torch.float32):
        raise ValueError(f"v must be float16 / bfloat16 / float32, got {v.dtype}")
    dev = v.device
    v = v.contiguous()
    s = s.to(torch.float32).contiguous()
    seq_len, num_states, num_heads, head_dim = v.shape
    v_merged = torch.empty(seq_len, num_heads, head_dim, dtype=v.dtype, device=dev)
    s_merged = torch.empty(seq_len, num_heads, dtype=torch.float32, device=dev)
    _lib.check(
        _lib.load().fi_merge_states(
            v.data_ptr(), s.data_ptr(), v_merged.data_ptr(), s_merged.data_ptr(),
            seq_len, num_states, num_heads, head_dim, dtype_code(v.dtype),
            device_index(dev), current_stream_ptr(dev),
        )
    )
    return v_merged, s_merged


def variable_length_merge_states(
    v: torch.Tensor, s: torch.Tensor, indptr: torch.Tensor,
    out_dtype: Optional[torch.dtype] = None,
) -> Tuple[torch.Tensor, torch.Tensor]:
    r"""Merge rows ``indptr[i]:indptr[i+1]`` of v ``[total, H, D]`` / s ``[total, H]`` into
    row i (reference VariableLengthMergeStates, cascade.cuh:686-736; empty ranges give
    v = 0, s = -5e4, cascade.cuh:397-405)."""
    if v.dim() != 3 or s.shape != v.shape[:2]:
        raise ValueError("v must be [total, H, D] and s [total, H]")
    if indptr.dtype != torch.int32:
        raise ValueError("indptr must be int32")
    dev = v.device
    v = v.contiguous()
    s = s.to(torch.float32).contiguous()
    indptr = indptr.to(dev).contiguous()
    out_dtype = out_dtype or v.dtype
    seq_len = indptr.numel() - 1
    _, num_heads, head_dim = v.shape
    v_merged = torch.empty(seq_len, num_heads, head_dim, dtype=out_dtype, device=dev)
    s_merged = torch.empty(seq_len, num_heads, dtype=torch.float32, device=dev)
    _lib.check(
        _lib.load().fi_variable_length_merge_states(
            v.data_ptr(), s.data_ptr(), indptr.data_ptr(), v_merged.data_ptr(),
            s_merged.data_ptr(), seq_len, num_heads, head_dim, dtype_code(v.dtype),
            dtype_code(out_dtype), 0, 1.0, device_index(dev), current_stream_ptr(dev),
        )
    )
    return v_merged, s_merged


# --------------------------------------------------------------------------- cascade wrappers
# the direct callers of decode + merge_state_in_place.  The reference
# builds them on its prefill kernels (cascade.py:228-556 on BatchPrefillWithPagedKVCacheWrapper,
# :558-796 on single_prefill_with_kv_cache).  This repo has no prefill path; a query row that
# attends a page list, with or without the causal mask, IS a decode request over a (possibly
# truncated) copy of that page list, so every level runs on the decode kernels: row t of an
# n-row group sees the group's first kv_len - (n - 1 - t) tokens under the causal mask
# (reference prefill.cuh mask: kv_idx <= kv_len - qo_len + qo_idx) and all kv_len otherwise.
# Rows of one group read the same pages: the duplicate reads are served by the 126 MB L2.


def rows_as_decode_requests(qo_indptr, kv_indptr, kv_indices, kv_last_page_len, page_size: int,
                            causal: bool):
    """Expand a (qo_indptr, page table) level into one decode request per query row.

    Returns ``(indptr, indices, last_page_len)`` of the per-row page table: int32 CPU tensors
    for indptr / last_page_len, ``indices`` gathered on ``kv_indices``' device.
    """
    import numpy as np

    qo = qo_indptr.detach().cpu().numpy().astype(np.int64)
    kvp = kv_indptr.detach().cpu().numpy().astype(np.int64)
    lpl = kv_last_page_len.detach().cpu().numpy().astype(np.int64)
    groups = len(qo) - 1
    if len(kvp) != groups + 1 or len(lpl) != groups:
        raise ValueError("qo_indptr, paged_kv_indptr and paged_kv_last_page_len disagree on the batch size")
    if groups < 0 or (groups >= 0 and qo[0] != 0) or np.any(np.diff(qo) < 0) or np.any(np.diff(kvp) < 0):
        raise ValueError("indptr arrays must start at 0 and be non-decreasing")
    rows = np.diff(qo)
    pages = np.diff(kvp)
    kv_len = np.where(pages > 0, (pages - 1) * page_size + lpl, 0)
    total = int(qo[-1]) if groups > 0 else 0
    grp = np.repeat(np.arange(groups), rows)
    t = np.arange(total) - qo[grp]
    row_len = kv_len[grp] - (rows[grp] - 1 - t) if causal else kv_len[grp]
    row_len = np.maximum(row_len, 0)
    row_pages = -(-row_len // page_size)
    row_indptr = np.zeros(total + 1, dtype=np.int64)
    np.cumsum(row_pages, out=row_indptr[1:])
    row_last = np.where(row_len > 0, (row_len - 1) % page_size + 1, 0)
    src = (np.arange(int(row_indptr[-1])) - np.repeat(row_indptr[:-1], row_pages) +
           np.repeat(kvp[grp], row_pages))
    gather = torch.from_numpy(src).to(kv_indices.device)
    row_indices = kv_indices.index_select(0, gather) if len(src) else kv_indices[:0]
    return (torch.from_numpy(row_indptr.astype(np.int32)), row_indices.to(torch.int32).contiguous(),
            torch.from_numpy(row_last.astype(np.int32)))


class MultiLevelCascadeAttentionWrapper:
    r"""Multi-level cascade attention over one unified page table (reference cascade.py:228-556).

    Same constructor, ``plan`` / ``run`` contract as the reference; every level is executed by a
    :class:`BatchDecodeWithPagedKVCacheWrapper` with one decode request per query row (see the
    note above), and the levels are combined with :func:`merge_state_in_place` in the
    reference's order (last level first, cascade.py:534-543).

    ``use_cuda_graph=True`` (reference cascade.py:313-330 takes per-level fixed buffers): ``run`` can
    be captured, ``plan`` stays host work between replays.  The caller's ``paged_kv_indices_buf_arr``
    gives the page capacity of every level; the per-ROW tables live in buffers of this wrapper
    (rows x that capacity entries, created by the first ``plan``), and the total number of query
    rows is fixed by that first ``plan`` (a captured graph holds the launch shapes).
    """

    def __init__(
        self,
        num_levels,
        float_workspace_buffer: torch.Tensor,
        kv_layout: str = "NHD",
        use_cuda_graph: bool = False,
        qo_indptr_buf_arr=None,
        paged_kv_indptr_buf_arr=None,
        paged_kv_indices_buf_arr=None,
        paged_kv_last_page_len_buf_arr=None,
    ) -> None:
        from .decode import BatchDecodeWithPagedKVCacheWrapper

        self._use_cuda_graph = bool(use_cuda_graph)
        self._float_workspace_buffer = float_workspace_buffer
        self._page_capacity = None
        if use_cuda_graph:
            if paged_kv_indices_buf_arr is None or len(paged_kv_indices_buf_arr) != num_levels:
                raise ValueError("use_cuda_graph needs paged_kv_indices_buf_arr with one buffer per level")
            self._page_capacity = [int(b.numel()) for b in paged_kv_indices_buf_arr]
            self._batch_decode_wrappers = [None] * num_levels  # fixed-size wrappers, made by the first plan()
        else:
            self._batch_decode_wrappers = [
                BatchDecodeWithPagedKVCacheWrapper(float_workspace_buffer, kv_layout)
                for _ in range(num_levels)
            ]
        self._num_levels = num_levels
        self._kv_layout = kv_layout
        self._num_rows = None

    @property
    def is_cuda_graph_enabled(self) -> bool:
        return self._use_cuda_graph

    def reset_workspace_buffer(self, float_workspace_buffer: torch.Tensor,
                               int_workspace_buffers) -> None:
        self._float_workspace_buffer = float_workspace_buffer
        for wrapper, int_ws in zip(self._batch_decode_wrappers, int_workspace_buffers):
            if wrapper is not None:
                wrapper.reset_workspace_buffer(float_workspace_buffer, int_ws)

    def plan(
        self,
        qo_indptr_arr,
        paged_kv_indptr_arr,
        paged_kv_indices_arr,
        paged_kv_last_page_len,
        num_qo_heads: int,
        num_kv_heads: int,
        head_dim: int,
        page_size: int,
        causal: bool = False,
        pos_encoding_mode: str = "NONE",
        use_fp16_qk_reduction: bool = False,
        sm_scale: Optional[float] = None,
        window_left: int = -1,
        logits_soft_cap: Optional[float] = None,
        rope_scale: Optional[float] = None,
        rope_theta: Optional[float] = None,
        q_data_type="float16",
        kv_data_type=None,
    ) -> None:
        if pos_encoding_mode != "NONE":
            # positions inside a level are level-local: the reference's own result for RoPE /
            # ALiBi across levels is not the un-cascaded attention either
            raise NotImplementedError("cascade attention supports pos_encoding_mode='NONE' only")
        if not (len(qo_indptr_arr) == len(paged_kv_indptr_arr) == len(paged_kv_indices_arr) ==
                len(paged_kv_last_page_len) == self._num_levels):
            raise ValueError("every per-level array must have num_levels entries")
        num_rows = None
        for i, (wrapper, qo, kvp, kvi, lpl) in enumerate(zip(
                self._batch_decode_wrappers, qo_indptr_arr, paged_kv_indptr_arr,
                paged_kv_indices_arr, paged_kv_last_page_len)):
            level_causal = causal if i == self._num_levels - 1 else False
            rows = self._uniform_rows(qo)
            kv_dt = canonicalize_torch_dtype(kv_data_type if kv_data_type is not None else q_data_type)
            level_rows = int(qo[-1])
            if num_rows is None:
                num_rows = level_rows
            elif num_rows != level_rows:
                raise ValueError("the last element of every qo_indptr must be the total number of queries")
            if (rows > 1 and not self._use_cuda_graph and window_left < 0
                    and rows_can_share_the_stream(kv_dt, head_dim, page_size)):
                # every group of the level has `rows` query rows: they share the group's KV stream as extra
                # columns of the GQA group (FiDecodeRunArgs.q_rows) instead of one decode request per row
                wrapper._q_rows, wrapper._q_rows_noncausal = rows, int(not level_causal)
                wrapper.plan(
                    kvp, kvi, lpl, rows * num_qo_heads, num_kv_heads, head_dim, page_size,
                    pos_encoding_mode="NONE", window_left=window_left, logits_soft_cap=logits_soft_cap,
                    q_data_type=q_data_type, kv_data_type=kv_data_type, sm_scale=sm_scale,
                    rope_scale=rope_scale, rope_theta=rope_theta,
                )
                continue
            indptr, indices, last = rows_as_decode_requests(qo, kvp, kvi, lpl, page_size, level_causal)
            if self._use_cuda_graph:
                wrapper = self._graph_level_wrapper(i, num_rows, indices)
            else:
                wrapper._q_rows, wrapper._q_rows_noncausal = 1, 0
            wrapper.plan(
                indptr, indices, last, num_qo_heads, num_kv_heads, head_dim, page_size,
                pos_encoding_mode="NONE", window_left=window_left, logits_soft_cap=logits_soft_cap,
                q_data_type=q_data_type, kv_data_type=kv_data_type, sm_scale=sm_scale,
                rope_scale=rope_scale, rope_theta=rope_theta,
            )
        self._num_rows = num_rows

    begin_forward = plan

    @staticmethod
    def _uniform_rows(qo_indptr) -> int:
        """Query rows per group when every group of the level has the same number (else 0)."""
        qo = qo_indptr.detach().cpu()
        if qo.numel() < 2:
            return 0
        d = qo[1:] - qo[:-1]
        n = int(d[0])
        return n if n > 0 and bool((d == n).all()) else 0

    def _graph_level_wrapper(self, level: int, num_rows: int, indices: torch.Tensor):
        """CUDA-graph mode: level `level`'s decode wrapper with fixed buffers for `num_rows` per-row
        requests and rows x (the level's page capacity) table entries."""
        from .decode import BatchDecodeWithPagedKVCacheWrapper

        if self._num_rows is not None and num_rows != self._num_rows:
            raise ValueError(f"the number of query rows is fixed in CUDA-graph mode ({self._num_rows}, got {num_rows})")
        capacity = max(num_rows * self._page_capacity[level], 1)
        if indices.numel() > capacity:
            raise ValueError("a level's page table exceeds the capacity of its paged_kv_indices buffer")
        w = self._batch_decode_wrappers[level]
        if w is None:
            dev = self._float_workspace_buffer.device
            w = BatchDecodeWithPagedKVCacheWrapper(
                self._float_workspace_buffer, self._kv_layout, use_cuda_graph=True,
                paged_kv_indptr_buffer=torch.zeros(num_rows + 1, dtype=torch.int32, device=dev),
                paged_kv_indices_buffer=torch.zeros(capacity, dtype=torch.int32, device=dev),
                paged_kv_last_page_len_buffer=torch.zeros(num_rows, dtype=torch.int32, device=dev))
            self._batch_decode_wrappers[level] = w
        return w

    def run(self, q: torch.Tensor, paged_kv_cache):
        if self._num_rows is None:
            raise RuntimeError("plan() must be called before run()")
        out, lse = self._batch_decode_wrappers[-1].run(q, paged_kv_cache, return_lse=True)
        for wrapper in self._batch_decode_wrappers[:-1]:
            out_i, lse_i = wrapper.run(q, paged_kv_cache, return_lse=True)
            merge_state_in_place(out, lse, out_i, lse_i)
        return out

    forward = run


def _shared_prefix_pages(k_shared: torch.Tensor, v_shared: torch.Tensor, kv_layout: str):
    """View a contiguous shared-prefix K / V (``[L, Hkv, D]`` NHD or ``[Hkv, L, D]`` HND) as a
    ``(k_pages, v_pages)`` paged cache without copying: pages of the largest power of two <= 64
    that divides L (the TMA loader's case) or single-token pages for any other L."""
    if k_shared.dim() != 3 or k_shared.shape != v_shared.shape:
        raise ValueError("k_shared / v_shared must be 3-D tensors of the same shape")
    k_shared, v_shared = k_shared.contiguous(), v_shared.contiguous()
    length = k_shared.shape[0] if kv_layout == "NHD" else k_shared.shape[1]
    page = 64
    while page >= 8 and length % page:
        page //= 2
    if page < 8:
        page = 1
    pages = length // page

    def view(x):
        if kv_layout == "NHD":
            return x.view(pages, page, x.shape[1], x.shape[2])
        return x.view(x.shape[0], pages, page, x.shape[2]).permute(1, 0, 2, 3)

    return view(k_shared), view(v_shared), pages, page


class BatchDecodeWithSharedPrefixPagedKVCacheWrapper:
    r"""Batch decode over a shared-prefix KV held in standalone tensors plus a per-request
    paged KV-cache (reference cascade.py:558-796; deprecated there in favour of
    :class:`MultiLevelCascadeAttentionWrapper`).

    The reference computes the shared level with ``single_prefill_with_kv_cache(causal=False)``;
    here it is a batch decode in which every request's page table is the same list of pages
    viewed over ``k_shared`` / ``v_shared`` in place (planned on first use per
    (prefix length, batch size) and cached).
    """

    def __init__(self, float_workspace_buffer: torch.Tensor, kv_layout: str = "NHD") -> None:
        from .decode import BatchDecodeWithPagedKVCacheWrapper

        self._batch_decode_wrapper = BatchDecodeWithPagedKVCacheWrapper(float_workspace_buffer, kv_layout)
        self._shared_wrapper = BatchDecodeWithPagedKVCacheWrapper(float_workspace_buffer, kv_layout)
        self._shared_key = None
        self._kv_layout = kv_layout

    def reset_workspace_buffer(self, float_workspace_buffer: torch.Tensor, int_workspace_buffer) -> None:
        self._batch_decode_wrapper.reset_workspace_buffer(float_workspace_buffer, int_workspace_buffer)
        self._shared_wrapper.reset_workspace_buffer(
            float_workspace_buffer, torch.empty_like(int_workspace_buffer))
        self._shared_key = None

    def begin_forward(
        self,
        unique_kv_indptr: torch.Tensor,
        unique_kv_indices: torch.Tensor,
        unique_kv_last_page_len: torch.Tensor,
        num_qo_heads: int,
        num_kv_heads: int,
        head_dim: int,
        page_size: int,
        data_type="float16",
    ) -> None:
        self._batch_decode_wrapper.begin_forward(
            unique_kv_indptr, unique_kv_indices, unique_kv_last_page_len, num_qo_heads,
            num_kv_heads, head_dim, page_size, pos_encoding_mode="NONE",
            q_data_type=data_type, kv_data_type=data_type,  # the reference passes data_type= only,
            # which leaves q at plan()'s float16 default: harmless there (fp16-only tests)
        )
        self._shape = (num_qo_heads, num_kv_heads, head_dim)
        self._shared_key = None

    def _plan_shared(self, batch_size: int, pages: int, page: int, q_dtype, kv_dtype) -> None:
        key = (batch_size, pages, page, q_dtype, kv_dtype)
        if key == self._shared_key:
            return
        num_qo_heads, num_kv_heads, head_dim = self._shape
        w = self._shared_wrapper
        if batch_size > 1 and rows_can_share_the_stream(kv_dtype, head_dim, page):
            # ONE request over the shared pages whose query rows are the batch (FiDecodeRunArgs.q_rows,
            # no mask among the rows): the prefix is streamed once per 8 (row, head) columns, not per request
            w._q_rows, w._q_rows_noncausal = batch_size, 1
            w.plan(torch.tensor([0, pages], dtype=torch.int32), torch.arange(pages, dtype=torch.int32),
                   torch.tensor([page if pages else 0], dtype=torch.int32), batch_size * num_qo_heads,
                   num_kv_heads, head_dim, page, pos_encoding_mode="NONE", q_data_type=q_dtype,
                   kv_data_type=kv_dtype, sm_scale=self._batch_decode_wrapper._sm_scale)
        else:
            w._q_rows, w._q_rows_noncausal = 1, 0
            indptr = torch.arange(batch_size + 1, dtype=torch.int32) * pages
            indices = torch.arange(pages, dtype=torch.int32).repeat(batch_size)
            last = torch.full((batch_size,), page if pages else 0, dtype=torch.int32)
            w.plan(
                indptr, indices, last, num_qo_heads, num_kv_heads, head_dim, page,
                pos_encoding_mode="NONE", q_data_type=q_dtype, kv_data_type=kv_dtype,
                sm_scale=self._batch_decode_wrapper._sm_scale,
            )
        self._shared_key = key

    def forward(self, q: torch.Tensor, k_shared: torch.Tensor, v_shared: torch.Tensor,
                unique_kv_cache) -> torch.Tensor:
        k_pages, v_pages, pages, page = _shared_prefix_pages(k_shared, v_shared, self._kv_layout)
        self._plan_shared(q.shape[0], pages, page, q.dtype, k_shared.dtype)
        V_shared, S_shared = self._shared_wrapper.run(q, (k_pages, v_pages), return_lse=True)
        V_unique, S_unique = self._batch_decode_wrapper.forward_return_lse(
            q, unique_kv_cache, pos_encoding_mode="NONE")
        merge_state_in_place(V_shared, S_shared, V_unique, S_unique)
        return V_shared

    def end_forward(self) -> None:
        r"""Deprecated in the reference; no effect."""
        pass
