"""Host-side helpers of the decode path (same names and behaviour as the small subset of
the reference's flashinfer/utils.py the path uses: enums :28-45, cache unpacking
:53-170, cached buffers :183-207, dtype canonicalisation :221-230, dtype / shape checks
:239-250, :493-511, PDL probe :556)."""
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


log2e = 1.44269504088896340736


def _check_pos_encoding_mode(pos_encoding_mode: str) -> None:
    if not hasattr(PosEncodingMode, pos_encoding_mode):
        raise KeyError("Invalid pos_encoding_mode {}".format(pos_encoding_mode))


def _check_kv_layout(kv_layout: str) -> None:
    if not hasattr(TensorLayout, kv_layout):
        raise KeyError("Invalid kv_layout {}".format(kv_layout))


def _expand_5d(x: torch.Tensor, kv_layout: str) -> torch.Tensor:
    if x.ndim not in [4, 5]:
        raise ValueError("x must be 4D or 5D")
    if x.ndim == 4:  # page_size == 1
        if kv_layout == "NHD":
            return x.unsqueeze(-3)
        elif kv_layout == "HND":
            return x.unsqueeze(-2)
        raise KeyError("Invalid kv_layout {}".format(kv_layout))
    return x


def _expand_4d(x: torch.Tensor, kv_layout: str) -> torch.Tensor:
    if x.ndim not in [3, 4]:
        raise ValueError("x must be 3D or 4D")
    if x.ndim == 3:  # page_size == 1
        if kv_layout == "NHD":
            return x.unsqueeze(-3)
        elif kv_layout == "HND":
            return x.unsqueeze(-2)
        raise KeyError("Invalid kv_layout {}".format(kv_layout))
    return x


def _unpack_paged_kv_cache(
    paged_kv_cache: Union[torch.Tensor, Tuple[torch.Tensor, torch.Tensor]],
    kv_layout: str,
) -> Tuple[torch.Tensor, torch.Tensor]:
    """(k_cache, v_cache) 4-D views; a 5-D cache is split on dim 1 (utils.py:149-170)."""
    if isinstance(paged_kv_cache, tuple):
        paged_k_cache, paged_v_cache = paged_kv_cache
        return _expand_4d(paged_k_cache, kv_layout), _expand_4d(paged_v_cache, kv_layout)
    elif torch.is_tensor(paged_kv_cache):
        paged_kv_cache = _expand_5d(paged_kv_cache, kv_layout)
        paged_k_cache, paged_v_cache = paged_kv_cache.unbind(dim=1)
        return paged_k_cache, paged_v_cache
    raise KeyError(
        "Unrecognized paged_kv_cache type {}, expect a single tensor or a tuple of tensor.".format(
            type(paged_kv_cache)
        )
    )


_cache_buf: Dict[Tuple[str, torch.device], torch.Tensor] = {}


def _get_cache_buf(name: str, bytes: int, device: torch.device) -> torch.Tensor:
    key = (name, device)
    buf = _cache_buf.get(key)
    if buf is None or buf.numel() < bytes:
        buf = torch.empty(bytes, dtype=torch.uint8, device=device)
        _cache_buf[key] = buf
    return buf


def get_alibi_slopes(n_heads: int) -> torch.Tensor:
    """ALiBi head slopes (reference flashinfer/utils.py:172-180): 2^(-8 i / n) for the largest
    power of two n <= n_heads, interleaved 2^(-4 (2j - 1) / n) for the remaining heads."""
    n = 2 ** math.floor(math.log2(n_heads))
    m_0 = 2.0 ** (-8.0 / n)
    m = torch.pow(m_0, torch.arange(1, 1 + n))
    if n < n_heads:
        m_hat_0 = 2.0 ** (-4.0 / n)
        m_hat = torch.pow(m_hat_0, torch.arange(1, 1 + 2 * (n_heads - n), 2))
        m = torch.cat([m, m_hat])
    return m.float()


def _get_cache_alibi_slopes_buf(num_qo_heads: int, device: torch.device) -> torch.Tensor:
    key = (f"alibi_slopes_{num_qo_heads}", device)
    buf = _cache_buf.get(key)
    if buf is None:
        buf = get_alibi_slopes(num_qo_heads).to(device)
        _cache_buf[key] = buf
    return buf


def _ceil_pow2(x: int) -> int:
    return 1 << (x - 1).bit_length()


def _get_range_buf(seq_len: int, device) -> torch.Tensor:
    seq_len_pow2 = _ceil_pow2(max(seq_len, 1))
    key = (f"range_{seq_len_pow2}", torch.device(device))
    buf = _cache_buf.get(key)
    if buf is None:
        buf = torch.arange(seq_len_pow2, device=device, dtype=torch.int32)
        _cache_buf[key] = buf
    return buf[:seq_len]


def canonicalize_torch_dtype(dtype: Union[torch.dtype, str]) -> torch.dtype:
    if isinstance(dtype, str):
        return getattr(torch, dtype)
    elif isinstance(dtype, torch.dtype):
        return dtype
    raise TypeError("dtype must be a string or torch.dtype, got {}".format(type(dtype)))


def _check_cached_qkv_data_type(
    q: torch.Tensor, k: torch.Tensor, dtype_q: torch.dtype, dtype_kv: torch.dtype
) -> None:
    if q.dtype != dtype_q:
        raise ValueError(
            f"The dtype of q {q.dtype} does not match the q_data_type {dtype_q} specified in plan function."
        )
    if k.dtype != dtype_kv:
        raise ValueError(
            f"The dtype of k {k.dtype} does not match the kv_data_type {dtype_kv} specified in plan function."
        )


def check_shape_dtype_device(
    x: torch.Tensor,
    expected_shape: Optional[Sequence[int]],
    expected_dtype: Optional[torch.dtype],
    expected_device: Optional[torch.device],
    name: str,
) -> None:
    if expected_shape and x.shape != torch.Size(expected_shape):
        raise ValueError(f"Invalid shape of {name}: expected {expected_shape}, got {x.shape}")
    if expected_dtype and x.dtype != expected_dtype:
        raise ValueError(f"Invalid dtype of {name}: expected {expected_dtype}, got {x.dtype}")
    if expected_device and x.device != expected_device:
        raise ValueError(f"Invalid device of {name}: expected {expected_device}, got {x.device}")


@functools.cache
@functools.lru_cache(maxsize=None)
def device_support_pdl(device: torch.device) -> bool:
    if device.type != "cuda":
        return False
    major, _ = torch.cuda.get_device_capability(device.index)
    return major >= 9


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
