"""Host-side helpers of the decode path.

Interface names that callers of the reference know (`PosEncodingMode`, `TensorLayout`,
`_unpack_paged_kv_cache`, `canonicalize_torch_dtype`, `check_shape_dtype_device`,
`get_alibi_slopes`, reference flashinfer/utils.py:28-45,149-180,221-250,493-511) keep their names,
argument meaning and exception classes; the implementations are this repo's own, and everything
the C-ABI path does not need was left out.
"""
from __future__ import annotations

import functools
import math
from enum import Enum
from typing import Dict, Optional, Sequence, Tuple, Union

import torch

from . import _lib


class PosEncodingMode(Enum):
    NONE = 0
    ROPE_LLAMA = 1
    ALIBI = 2


class TensorLayout(Enum):
    NHD = 0
    HND = 1


_POS_MODES = frozenset(m.name for m in PosEncodingMode)
_LAYOUTS = frozenset(m.name for m in TensorLayout)


def _check_pos_encoding_mode(pos_encoding_mode: str) -> None:
    if pos_encoding_mode not in _POS_MODES:
        raise KeyError(f"Invalid pos_encoding_mode {pos_encoding_mode}")


def _check_kv_layout(kv_layout: str) -> None:
    if kv_layout not in _LAYOUTS:
        raise KeyError(f"Invalid kv_layout {kv_layout}")


def _with_page_axis(x: torch.Tensor, full_ndim: int, kv_layout: str) -> torch.Tensor:
    """A cache with `page_size == 1` may come without its page axis (one dimension short of
    `full_ndim`): put the axis of length one where the layout has it."""
    if x.ndim == full_ndim:
        return x
    if x.ndim != full_ndim - 1:
        raise ValueError(f"x must be {full_ndim - 1}D or {full_ndim}D")
    _check_kv_layout(kv_layout)
    return x.unsqueeze(-3 if kv_layout == "NHD" else -2)


def _unpack_paged_kv_cache(
    paged_kv_cache: Union[torch.Tensor, Tuple[torch.Tensor, torch.Tensor]],
    kv_layout: str,
) -> Tuple[torch.Tensor, torch.Tensor]:
    """(k_cache, v_cache) as 4-D views `[pages, page_size, Hkv, D]` (NHD) / `[pages, Hkv, page_size,
    D]` (HND): a tuple passes through, a single tensor holds K and V as the two halves of dim 1."""
    if isinstance(paged_kv_cache, tuple):
        k_cache, v_cache = paged_kv_cache
        return _with_page_axis(k_cache, 4, kv_layout), _with_page_axis(v_cache, 4, kv_layout)
    if torch.is_tensor(paged_kv_cache):
        both = _with_page_axis(paged_kv_cache, 5, kv_layout)
        return both[:, 0], both[:, 1]
    raise KeyError(
        f"Unrecognized paged_kv_cache type {type(paged_kv_cache)}, expect a single tensor or a tuple of tensor."
    )


# scratch tensors that outlive a call, one per (purpose, device)
_buffers: Dict[Tuple[str, torch.device], torch.Tensor] = {}


def _get_cache_buf(name: str, bytes: int, device: torch.device) -> torch.Tensor:
    """A uint8 scratch tensor of at least `bytes` bytes, grown on demand, never zeroed."""
    buf = _buffers.get((name, device))
    if buf is None or buf.numel() < bytes:
        buf = _buffers[(name, device)] = torch.empty(bytes, dtype=torch.uint8, device=device)
    return buf


def get_alibi_slopes(n_heads: int) -> torch.Tensor:
    """ALiBi head slopes (Press et al.; the reference's values, flashinfer/utils.py:172-180): with n the
    largest power of two <= n_heads, head i < n gets 2^(-8 (i + 1) / n); the remaining heads take
    every other slope of the 2n-head series, 2^(-4 (2 j + 1) / n)."""
    n = 1 << (max(n_heads, 1).bit_length() - 1)
    # evaluated in float32 by torch.pow like the reference, so the slopes match it bit for bit
    slopes = torch.pow(2.0 ** (-8.0 / n), torch.arange(1, n + 1))
    if n_heads > n:
        extra = torch.pow(2.0 ** (-4.0 / n), torch.arange(1, 2 * (n_heads - n), 2))
        slopes = torch.cat([slopes, extra])
    return slopes.float()


def _get_cache_alibi_slopes_buf(num_qo_heads: int, device: torch.device) -> torch.Tensor:
    key = (f"alibi_slopes_{num_qo_heads}", device)
    if key not in _buffers:
        _buffers[key] = get_alibi_slopes(num_qo_heads).to(device)
    return _buffers[key]


def canonicalize_torch_dtype(dtype: Union[torch.dtype, str]) -> torch.dtype:
    if isinstance(dtype, torch.dtype):
        return dtype
    if isinstance(dtype, str):
        return getattr(torch, dtype)
    raise TypeError(f"dtype must be a string or torch.dtype, got {type(dtype)}")


def check_shape_dtype_device(
    x: torch.Tensor,
    expected_shape: Optional[Sequence[int]],
    expected_dtype: Optional[torch.dtype],
    expected_device: Optional[torch.device],
    name: str,
) -> None:
    if expected_shape and tuple(x.shape) != tuple(expected_shape):
        raise ValueError(f"Invalid shape of {name}: expected {expected_shape}, got {x.shape}")
    if expected_dtype and x.dtype != expected_dtype:
        raise ValueError(f"Invalid dtype of {name}: expected {expected_dtype}, got {x.dtype}")
    if expected_device and x.device != expected_device:
        raise ValueError(f"Invalid device of {name}: expected {expected_device}, got {x.device}")


@functools.lru_cache(maxsize=None)
def device_support_pdl(device: torch.device) -> bool:
    """Programmatic dependent launch needs sm_90 or later."""
    return device.type == "cuda" and torch.cuda.get_device_capability(device.index)[0] >= 9


def is_float8(x: torch.Tensor) -> bool:
    return x.dtype in (torch.float8_e4m3fn, torch.float8_e5m2)


_DTYPE_CODE = {
    torch.float16: _lib.DTYPE_F16,
    torch.bfloat16: _lib.DTYPE_BF16,
    torch.float8_e4m3fn: _lib.DTYPE_E4M3,
    torch.float8_e5m2: _lib.DTYPE_E5M2,
    torch.float32: _lib.DTYPE_F32,
}


def dtype_code(dtype: torch.dtype) -> int:
    """Map a torch dtype to the C-ABI's FI_DTYPE_* code (include/fi_b200.h)."""
    try:
        return _DTYPE_CODE[dtype]
    except KeyError:
        raise ValueError(f"unsupported dtype {dtype}") from None


def device_index(device: torch.device) -> int:
    if device.type != "cuda":
        raise RuntimeError(
            f"flashinfer_b200 ops run on CUDA tensors only (got device {device}); there is no CPU fallback"
        )
    return device.index if device.index is not None else torch.cuda.current_device()


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def current_stream_ptr(device: torch.device) -> int:
    """cudaStream_t of torch's current stream on `device` (the raw getter skips the Stream object:
    ~0.3 us instead of ~2 us on every plan() / run())."""
    if _raw_stream is not None and device.index is not None:
        return _raw_stream(device.index)
    return torch.cuda.current_stream(device).cuda_stream


def default_sm_scale(head_dim: int) -> float:
    return 1.0 / math.sqrt(head_dim)
