"""Page-table maintenance around the decode path (reference flashinfer/page.py:169-415):
``get_seq_lens``, ``get_batch_indices_positions`` and ``append_paged_kv_cache`` — the step that
writes the new tokens' K / V into the paged cache right before every decode call.

Same names, arguments and error behaviour as the reference; the device work is done by the
hand-written kernels in csrc/page.cu through the C-ABI (the reference uses a Triton kernel for
``get_batch_indices_positions`` and a JIT-compiled CUDA kernel for the append).  No CPU fallback.
"""
from __future__ import annotations

from typing import Tuple, Union

import torch

from . import _lib
from .utils import (
    _check_kv_layout,
    _unpack_paged_kv_cache,
    current_stream_ptr,
    device_index,
    dtype_code,
)


def get_seq_lens(
    kv_indptr: torch.Tensor, kv_last_page_len: torch.Tensor, page_size: int
) -> torch.Tensor:
    r"""Convert KV indptr and last page length to sequence lengths.

    ``kv_len[i] = max(indptr[i+1] - indptr[i] - 1, 0) * page_size + last_page_len[i]``
    (reference flashinfer/page.py:243-246).
    """
    return (
        torch.clamp(kv_indptr[1:] - kv_indptr[:-1] - 1, min=0) * page_size
        + kv_last_page_len
    )


def _i32_on(t: torch.Tensor, device: torch.device, name: str) -> torch.Tensor:
    if t.device != device:
        raise ValueError(f"{name} must be on {device}, got {t.device}")
    if t.dtype != torch.int32:
        t = t.to(torch.int32)  # the reference's Triton kernel loads whatever dtype it is given
    return t.contiguous()


def get_batch_indices_positions(
    append_indptr: torch.Tensor, seq_lens: torch.Tensor, nnz: int
) -> Tuple[torch.Tensor, torch.Tensor]:
    r"""Convert append indptr and sequence lengths to batch indices and positions
    (reference flashinfer/page.py:169-221): for ``append_indptr[i] <= j < append_indptr[i+1]``,
    ``batch_indices[j] = i`` and ``positions[j] = j + seq_lens[i] - append_indptr[i+1]``.

    ``seq_lens`` may live on another device (the reference example passes a CPU tensor); it is
    moved next to ``append_indptr``.  Returns two int32 tensors of ``nnz`` entries.
    """
    dev = append_indptr.device
    dev_idx = device_index(dev)
    batch_size = append_indptr.size(0) - 1
    append_indptr = _i32_on(append_indptr, dev, "append_indptr")
    seq_lens = _i32_on(seq_lens.to(dev), dev, "seq_lens")
    if seq_lens.numel() != batch_size:
        raise ValueError("seq_lens must have batch_size entries")
    batch_indices = torch.empty((nnz,), device=dev, dtype=torch.int32)
    positions = torch.empty((nnz,), device=dev, dtype=torch.int32)
    lib = _lib.load()
    _lib.check(
        lib.fi_get_batch_indices_positions(
            append_indptr.data_ptr(), seq_lens.data_ptr(), batch_indices.data_ptr(),
            positions.data_ptr(), int(batch_size), int(nnz), dev_idx, current_stream_ptr(dev),
        )
    )
    return batch_indices, positions


def append_paged_kv_cache(
    append_key: torch.Tensor,
    append_value: torch.Tensor,
    batch_indices: torch.Tensor,
    positions: torch.Tensor,
    paged_kv_cache: Union[torch.Tensor, Tuple[torch.Tensor, torch.Tensor]],
    kv_indices: torch.Tensor,
    kv_indptr: torch.Tensor,
    kv_last_page_len: torch.Tensor,
    kv_layout: str = "NHD",
) -> None:
    r"""Append a batch of key-value pairs to a paged key-value cache
    (reference flashinfer/page.py:299-415).

    append_key / append_value ``[nnz, num_kv_heads, head_dim]`` (views with unit stride on
    head_dim are fine), batch_indices / positions ``[nnz]`` from
    :func:`get_batch_indices_positions`, paged_kv_cache a 5-D tensor or a ``(k, v)`` tuple in
    ``kv_layout``.  The pages must already be allocated: ``kv_indices`` / ``kv_indptr`` /
    ``kv_last_page_len`` describe the cache *after* the append.  In place, returns ``None``.
    """
    _check_kv_layout(kv_layout)
    k_cache, v_cache = _unpack_paged_kv_cache(paged_kv_cache, kv_layout)
    dev = append_key.device
    dev_idx = device_index(dev)
    if append_key.dim() != 3 or append_value.shape != append_key.shape:
        raise ValueError("append_key / append_value must both have shape [nnz, num_kv_heads, head_dim]")
    nnz, num_kv_heads, head_dim = append_key.shape
    if append_value.dtype != append_key.dtype or k_cache.dtype != append_key.dtype \
            or v_cache.dtype != append_key.dtype:
        raise ValueError("append_key, append_value and the paged cache must share one dtype")
    for t, name in [(append_value, "append_value"), (k_cache, "k_cache"), (v_cache, "v_cache"),
                    (batch_indices, "batch_indices"), (positions, "positions"),
                    (kv_indices, "kv_indices"), (kv_indptr, "kv_indptr")]:
        if t.device != dev:
            raise ValueError(f"{name} must be on {dev}, got {t.device}")
    if batch_indices.numel() != nnz or positions.numel() != nnz:
        raise ValueError("batch_indices / positions must have nnz entries")
    if kv_layout == "NHD":
        page_size, cache_heads = k_cache.shape[1], k_cache.shape[2]
        stride_n, stride_h = k_cache.stride(1), k_cache.stride(2)
    else:
        cache_heads, page_size = k_cache.shape[1], k_cache.shape[2]
        stride_h, stride_n = k_cache.stride(1), k_cache.stride(2)
    if cache_heads != num_kv_heads or k_cache.shape[3] != head_dim:
        raise ValueError("paged_kv_cache does not match the (num_kv_heads, head_dim) of append_key")
    if k_cache.stride() != v_cache.stride():
        raise ValueError("k/v strides must be identical")
    if append_key.stride(-1) != 1 or append_value.stride(-1) != 1 or k_cache.stride(-1) != 1:
        raise ValueError("the last dimension of append_key / append_value / the cache must be contiguous")
    batch_indices = _i32_on(batch_indices, dev, "batch_indices")
    positions = _i32_on(positions, dev, "positions")
    kv_indices = _i32_on(kv_indices, dev, "kv_indices")
    kv_indptr = _i32_on(kv_indptr, dev, "kv_indptr")
    lib = _lib.load()
    _lib.check(
        lib.fi_append_paged_kv_cache(
            append_key.data_ptr(), append_value.data_ptr(),
            append_key.stride(0), append_key.stride(1), append_value.stride(0), append_value.stride(1),
            batch_indices.data_ptr(), positions.data_ptr(), int(nnz),
            k_cache.data_ptr(), v_cache.data_ptr(), k_cache.stride(0), stride_n, stride_h,
            kv_indices.data_ptr(), kv_indptr.data_ptr(), int(num_kv_heads), int(head_dim),
            int(page_size), dtype_code(append_key.dtype), dev_idx, current_stream_ptr(dev),
        )
    )
