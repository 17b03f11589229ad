"""Attention-state merge operators — same names and semantics as the reference's
flashinfer/cascade.py (merge_state :46-99, merge_state_in_place :112-158,
merge_states :172-215), backed by csrc/merge.cu through the C-ABI.

A state is (v, s): v the normalised partial attention output, s its log-sum-exp in
**base 2** (reference include/flashinfer/attention/state.cuh:45, cascade.cuh:55-69).
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import _lib
from .utils import current_stream_ptr, device_index, dtype_code


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
