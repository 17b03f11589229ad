"""Decode attention API — same public names, signatures and error behaviour as the
reference's flashinfer/decode.py (single_decode_with_kv_cache :389-578,
BatchDecodeWithPagedKVCacheWrapper :581-1410, CUDAGraphBatchDecodeWithPagedKVCacheWrapper
:1413-1480, fast_decode_plan :2416-2579), backed by the prebuilt sm_100a C-ABI library
(include/fi_b200.h) instead of per-configuration JIT modules.

Host code is Python/PyTorch (device memory, streams); all arithmetic happens in the CUDA
kernels behind libfi_b200.so.  There is no CPU path: tensors must live on a CUDA device.
"""
from __future__ import annotations

import ctypes as C
import functools
import math
import os
from typing import Any, List, Optional, Tuple, Union

import torch

from . import _lib
from .page import get_seq_lens
from .utils import (
    _check_kv_layout,
    _check_pos_encoding_mode,
    PosEncodingMode,
    _get_cache_alibi_slopes_buf,
    _get_cache_buf,
    _unpack_paged_kv_cache,
    canonicalize_torch_dtype,
    check_shape_dtype_device,
    current_stream_ptr,
    device_index,
    device_support_pdl,
    dtype_code,
    is_float8,
    TensorLayout,
)

_INT_WORKSPACE_BYTES = 8 * 1024 * 1024
_PI_OFF_INDICES = 28  # PlanInfoSlot kPiOffIndices (csrc/decode_params.h)
_PlanInfo = C.c_int64 * 32  # FI_PLAN_INFO_LEN; a ctypes array: no tensor object on the per-step path


def _host_i32_ptr(t: torch.Tensor) -> Tuple[torch.Tensor, int]:
    """Contiguous CPU int32 copy of `t` and its address (kept alive by the caller)."""
    if t.device.type == "cpu" and t.dtype == torch.int32 and t.is_contiguous():
        return t, t.data_ptr()
    h = t.detach().to("cpu", torch.int32).contiguous()
    return h, h.data_ptr()


def rows_can_share_the_stream(kv_dtype: torch.dtype, head_dim: int, page_size: int) -> bool:
    """Shapes with multi-row units (FiDecodeRunArgs.q_rows): 16-bit KV on the mma.sync kernel, fp8 KV on the
    tcgen05 kernel's wide mode (head_dim 128, TMA-sized pages)."""
    if os.environ.get("FI_B200_MTP_ROWS", "1") == "0" or head_dim == 512:
        return False
    if kv_dtype in (torch.float16, torch.bfloat16):
        return True
    return (kv_dtype in (torch.float8_e4m3fn, torch.float8_e5m2) and head_dim == 128 and page_size >= 8
            and page_size & (page_size - 1) == 0 and os.environ.get("FI_B200_FP8_TC5", "1") != "0")


def _unsupported_variant(pos_encoding_mode: str) -> None:
    _check_pos_encoding_mode(pos_encoding_mode)
    # all three reference modes are built: "NONE" and "ALIBI" on the tuned kernels, "ROPE_LLAMA"
    # on the CUDA-core variant kernel (csrc/decode_rope_kernel.cuh)


def _pos_mode_args(pos_encoding_mode: str, num_qo_heads: int, device: torch.device,
                   rope_scale: Optional[float], rope_theta: Optional[float]):
    """(mode code, alibi slopes pointer, rope_scale, rope_theta) for the C-ABI run functions;
    defaults as the reference (decode.py:1257-1260: rope_scale 1.0, rope_theta 1e4)."""
    return (
        PosEncodingMode[pos_encoding_mode].value,
        _alibi_slopes_ptr(pos_encoding_mode, num_qo_heads, device),
        float(rope_scale) if rope_scale is not None else 1.0,
        float(rope_theta) if rope_theta is not None else 1e4,
    )


def _alibi_slopes_ptr(pos_encoding_mode: str, num_qo_heads: int, device: torch.device):
    """Device pointer of the cached ALiBi slopes (reference decode.py:530,1306:
    `_get_cache_alibi_slopes_buf`) when the mode is "ALIBI", else None."""
    if pos_encoding_mode != "ALIBI":
        return None
    return _get_cache_alibi_slopes_buf(num_qo_heads, device).data_ptr()


def single_decode_with_kv_cache(
    q: torch.Tensor,
    k: torch.Tensor,
    v: torch.Tensor,
    kv_layout: str = "NHD",
    pos_encoding_mode: str = "NONE",
    use_tensor_cores: bool = False,
    q_scale: Optional[float] = None,
    k_scale: Optional[float] = None,
    v_scale: Optional[float] = None,
    window_left: int = -1,
    logits_soft_cap: Optional[float] = None,
    sm_scale: Optional[float] = None,
    rope_scale: Optional[float] = None,
    rope_theta: Optional[float] = None,
    return_lse: bool = False,
) -> Union[torch.Tensor, Tuple[torch.Tensor, torch.Tensor]]:
    r"""Decode attention with KV cache for a single request (reference decode.py:389-578).

    q ``[num_qo_heads, head_dim]``; k, v ``[kv_len, num_kv_heads, head_dim]`` (NHD) or
    ``[num_kv_heads, kv_len, head_dim]`` (HND), contiguous (csrc/single_decode.cu:36-39).
    Returns o ``[num_qo_heads, head_dim]`` (and base-2 lse ``[num_qo_heads]``).
    ``use_tensor_cores`` is accepted and ignored: the sm_100a kernel always runs its
    MMA path.
    """
    _check_kv_layout(kv_layout)
    _unsupported_variant(pos_encoding_mode)
    if q.dim() != 2 or k.dim() != 3 or v.dim() != 3:
        raise ValueError("q must be 2-D and k / v 3-D")
    if k.shape != v.shape:
        raise ValueError("k and v must have the same shape")
    if k.dtype != v.dtype:
        raise ValueError("k and v must have the same dtype")
    for t, name in ((q, "q"), (k, "k"), (v, "v")):
        if not t.is_contiguous():
            raise ValueError(f"{name} must be contiguous")
    dev = q.device
    dev_idx = device_index(dev)
    if k.device != dev or v.device != dev:
        raise ValueError("q, k, v must be on the same device")
    num_qo_heads, head_dim = q.shape
    if kv_layout == "NHD":
        kv_len, num_kv_heads = k.shape[0], k.shape[1]
    else:
        num_kv_heads, kv_len = k.shape[0], k.shape[1]
    if k.shape[2] != head_dim:
        raise ValueError("head_dim of q and k/v must match")
    if logits_soft_cap is None:
        logits_soft_cap = 0.0
    if sm_scale is None:
        sm_scale = 1.0 / math.sqrt(head_dim)
    if q_scale is not None:
        sm_scale *= q_scale
    if k_scale is not None:
        sm_scale *= k_scale

    lib = _lib.load()
    tmp_bytes = lib.fi_single_decode_tmp_bytes(num_qo_heads, num_kv_heads, head_dim, dev_idx)
    tmp = _get_cache_buf("single_decode_with_kv_cache_tmp", max(tmp_bytes, 32 * 1024 * 1024), dev)
    o = torch.empty_like(q)
    lse = torch.empty((num_qo_heads,), dtype=torch.float32, device=dev) if return_lse else None
    _lib.check(
        lib.fi_single_decode_run(
            q.data_ptr(), k.data_ptr(), v.data_ptr(), tmp.data_ptr(), tmp.numel(),
            o.data_ptr(), lse.data_ptr() if lse is not None else None,
            num_qo_heads, num_kv_heads, head_dim, kv_len,
            dtype_code(q.dtype), dtype_code(k.dtype), TensorLayout[kv_layout].value,
            window_left, float(sm_scale), float(v_scale) if v_scale is not None else 1.0,
            float(logits_soft_cap),
            *_pos_mode_args(pos_encoding_mode, num_qo_heads, dev, rope_scale, rope_theta),
            dev_idx, current_stream_ptr(dev),
        )
    )
    return (o, lse) if return_lse else o


class BatchDecodeWithPagedKVCacheWrapper:
    r"""Batch decode over a paged KV cache (reference decode.py:581-1410).

    >>> wrapper = BatchDecodeWithPagedKVCacheWrapper(workspace, "NHD")
    >>> wrapper.plan(indptr, indices, last_page_len, Hq, Hkv, D, page_size,
    ...              q_data_type=torch.bfloat16, kv_data_type=torch.bfloat16)
    >>> o = wrapper.run(q, paged_kv_cache)

    ``plan`` runs the host scheduler (C++, csrc/plan.cu) once per decode step;
    ``run`` launches the persistent sm_100a kernel (+ the split-KV merge) per layer and
    never synchronises, so it can be captured in a CUDA graph.
    """

    def __init__(
        self,
        float_workspace_buffer: torch.Tensor,
        kv_layout: str = "NHD",
        use_cuda_graph: bool = False,
        use_tensor_cores: bool = False,
        paged_kv_indptr_buffer: Optional[torch.Tensor] = None,
        paged_kv_indices_buffer: Optional[torch.Tensor] = None,
        paged_kv_last_page_len_buffer: Optional[torch.Tensor] = None,
        backend: str = "auto",
        jit_args: Optional[List[Any]] = None,
    ) -> None:
        _check_kv_layout(kv_layout)
        if jit_args is not None:
            # custom JIT attention variants are out of scope (no JIT here)
            raise NotImplementedError("jit_args (custom JIT variants) are not supported")
        self._jit_module = None
        self._kv_layout = kv_layout
        self._float_workspace_buffer = float_workspace_buffer
        self.device = float_workspace_buffer.device
        device_index(self.device)  # raises on non-CUDA devices: no CPU fallback
        self._int_workspace_buffer = torch.empty(
            (_INT_WORKSPACE_BYTES,), dtype=torch.uint8, device=self.device
        )
        self._pin_memory_int_workspace_buffer = torch.empty(
            (_INT_WORKSPACE_BYTES,), dtype=torch.uint8, pin_memory=True, device="cpu"
        )
        if use_cuda_graph:
            if not torch.is_tensor(paged_kv_indptr_buffer):
                raise ValueError("paged_kv_indptr_buffer should be a torch.Tensor in cudagraph mode")
            if not torch.is_tensor(paged_kv_indices_buffer):
                raise ValueError("paged_kv_indices_buffer should be a torch.Tensor in cudagraph mode")
            if not torch.is_tensor(paged_kv_last_page_len_buffer):
                raise ValueError(
                    "paged_kv_last_page_len_buffer should be a torch.Tensor in cudagraph mode"
                )
            self._fixed_batch_size = len(paged_kv_last_page_len_buffer)
            if len(paged_kv_indptr_buffer) != self._fixed_batch_size + 1:
                raise ValueError("The size of paged_kv_indptr_buffer should be batch_size + 1")
        else:
            self._fixed_batch_size = 0
        self._indptr_dev = self._indptr_src = None
        self._last_page_len_dev = self._last_page_len_src = None
        self._indices_dev = self._indices_src = None
        self._paged_kv_indptr_buf = paged_kv_indptr_buffer
        self._paged_kv_indices_buf = paged_kv_indices_buffer
        self._paged_kv_last_page_len_buf = paged_kv_last_page_len_buffer
        # both flags select the same kernel here: the MMA path is the only path
        self._use_tensor_cores = use_tensor_cores or backend == "trtllm-gen"
        self._use_cuda_graph = use_cuda_graph
        self._backend = backend
        self._plan_info = None  # opaque (FI_PLAN_INFO_LEN x int64), written by the C planner
        # argument blocks of the C-ABI calls (include/fi_b200.h FiDecodePlanArgs / FiDecodeRunArgs):
        # filled once, per step only the fields that change are rewritten and the call takes ONE
        # pointer (the 26 per-call argument conversions were a third of plan()'s / run()'s host time)
        self._pa, self._ra = _lib.DecodePlanArgs(), _lib.DecodeRunArgs()
        self._pa_ref, self._ra_ref = C.byref(self._pa), C.byref(self._ra)
        self._bound_ws = (None, None, None)
        self._kv_key = None            # (weakref to the last 5-D cache tensor, its data_ptr)
        self._pos_key = None
        self._planned_window_left = -1
        self._kv_sig = None
        self._plan_sig = None
        self._dev_idx = device_index(self.device)
        self._lib = _lib.load()
        self._pdl_default = int(device_support_pdl(self.device))
        self._head_pairs_ok = True   # cleared when a run() had to fall back (FI_ERR_REPLAN)
        self._ran_once = False       # step(): the run argument block holds a complete run()
        self._q_rows = 1             # > 1: a multi-token child whose rows share the KV stream (q_rows)
        self._q_rows_noncausal = 0   # cascade levels below the last: no mask among the rows
        self._run_kw = (None, None, None, None, None)
        self._plan_call = None
        self._mtp_children = {}
        self._mtp_src = None
        self._pos_encoding_mode = "NONE"
        self._rope_scale = self._rope_theta = None

    @property
    def use_tensor_cores(self) -> bool:
        return self._use_tensor_cores

    # The kernels take every request's page base and length from the planned work list, so outside
    # CUDA-graph mode the device copies of indptr / last_page_len (attributes the reference keeps,
    # decode.py:1009-1017) are made on first access instead of on every plan().
    @property
    def _paged_kv_indptr_buf(self):
        if self._indptr_dev is None and self._indptr_src is not None:
            self._indptr_dev = self._indptr_src.to(self.device, non_blocking=True)
        return self._indptr_dev

    @_paged_kv_indptr_buf.setter
    def _paged_kv_indptr_buf(self, value):
        self._indptr_dev, self._indptr_src = value, None

    # The page table itself: when plan() gets it on the host, the C planner appends it to the work
    # list (one H2D copy for everything) and the kernels read that copy; the attribute the reference
    # keeps (decode.py:1009-1012) is materialised on first access.
    @property
    def _paged_kv_indices_buf(self):
        if self._indices_dev is None and self._indices_src is not None:
            self._indices_dev = self._indices_src.to(self.device, non_blocking=True)
        return self._indices_dev

    @_paged_kv_indices_buf.setter
    def _paged_kv_indices_buf(self, value):
        self._indices_dev, self._indices_src = value, None

    @property
    def _paged_kv_last_page_len_buf(self):
        if self._last_page_len_dev is None and self._last_page_len_src is not None:
            self._last_page_len_dev = self._last_page_len_src.to(self.device, non_blocking=True)
        return self._last_page_len_dev

    @_paged_kv_last_page_len_buf.setter
    def _paged_kv_last_page_len_buf(self, value):
        self._last_page_len_dev, self._last_page_len_src = value, None

    @property
    def is_cuda_graph_enabled(self) -> bool:
        return self._use_cuda_graph

    def reset_workspace_buffer(
        self, float_workspace_buffer: torch.Tensor, int_workspace_buffer: torch.Tensor
    ) -> None:
        self._float_workspace_buffer = float_workspace_buffer
        self._int_workspace_buffer = int_workspace_buffer
        self._pin_memory_int_workspace_buffer = torch.empty(
            self._int_workspace_buffer.shape,
            dtype=self._int_workspace_buffer.dtype,
            device="cpu",
            pin_memory=True,
        )

    # ------------------------------------------------------------------ plan
    def plan(
        self,
        indptr: torch.Tensor,
        indices: torch.Tensor,
        last_page_len: torch.Tensor,
        num_qo_heads: int,
        num_kv_heads: int,
        head_dim: int,
        page_size: int,
        pos_encoding_mode: str = "NONE",
        window_left: int = -1,
        logits_soft_cap: Optional[float] = None,
        q_data_type: Optional[Union[str, torch.dtype]] = "float16",
        kv_data_type: Optional[Union[str, torch.dtype]] = None,
        data_type: Optional[Union[str, torch.dtype]] = None,
        sm_scale: Optional[float] = None,
        rope_scale: Optional[float] = None,
        rope_theta: Optional[float] = None,
        non_blocking: bool = True,
        block_tables: Optional[torch.Tensor] = None,
        seq_lens: Optional[torch.Tensor] = None,
        fixed_split_size: Optional[int] = None,
        disable_split_kv: bool = False,
    ) -> None:
        r"""Plan batch decode for a page table (reference decode.py:810-1102).

        Cannot be captured in a CUDA graph (it reads indptr / last_page_len on the host).
        """
        i32 = torch.int32
        if indptr.dtype != i32 or indices.dtype != i32 or last_page_len.dtype != i32:
            for tensor, name in [(indptr, "indptr"), (indices, "indices"), (last_page_len, "last_page_len")]:
                if tensor.dtype != i32:
                    raise ValueError(f"{name} must have dtype torch.int32, got {tensor.dtype}")
        if pos_encoding_mode != "NONE":
            _unsupported_variant(pos_encoding_mode)
        self._workspace_size = (
            self._float_workspace_buffer.numel() * self._float_workspace_buffer.element_size()
        )
        batch_size = last_page_len.shape[0]
        if logits_soft_cap is None:
            logits_soft_cap = 0.0

        if self._use_cuda_graph:
            if batch_size != self._fixed_batch_size:
                raise ValueError(
                    "The batch size should be fixed in cudagraph mode, the runtime batch size {} "
                    " mismatches the batch size set during initialization {}".format(
                        batch_size, self._fixed_batch_size
                    )
                )
            if len(indices) > len(self._paged_kv_indices_buf):
                raise ValueError(
                    "The size of indices should be less than or equal to the allocated buffer"
                )
            self._paged_kv_indptr_buf.copy_(indptr, non_blocking=non_blocking)
            self._paged_kv_last_page_len_buf.copy_(last_page_len, non_blocking=non_blocking)
            self._paged_kv_indices_buf[: len(indices)].copy_(
                indices, non_blocking=(indices.device == self.device) and non_blocking
            )
        else:
            if indices.device.type == "cpu":
                self._indices_dev, self._indices_src = None, indices  # the planner copies it (below)
            else:
                self._paged_kv_indices_buf = indices.to(self.device, non_blocking=non_blocking)
            if indptr.device == self.device:
                self._paged_kv_indptr_buf = indptr
            else:
                self._indptr_dev, self._indptr_src = None, indptr
            if last_page_len.device == self.device:
                self._paged_kv_last_page_len_buf = last_page_len
            else:
                self._last_page_len_dev, self._last_page_len_src = None, last_page_len

        indptr_host, indptr_ptr = _host_i32_ptr(indptr)
        last_page_len_host, lpl_ptr = _host_i32_ptr(last_page_len)
        if indptr_host.shape[0] != batch_size + 1:
            raise ValueError("indptr must have batch_size + 1 entries")

        if data_type is not None:
            if q_data_type is None:
                q_data_type = data_type
            if kv_data_type is None:
                kv_data_type = data_type
        q_data_type = canonicalize_torch_dtype(q_data_type)
        if kv_data_type is None:
            kv_data_type = q_data_type
        kv_data_type = canonicalize_torch_dtype(kv_data_type)
        if fixed_split_size is not None and not self.use_tensor_cores:
            raise ValueError("fixed_split_size is only supported by tensor core decode for now.")

        self._cached_q_data_type = q_data_type
        self._cached_kv_data_type = kv_data_type
        self._batch_size = batch_size
        self._num_qo_heads = num_qo_heads
        self._num_kv_heads = num_kv_heads
        self._block_tables = block_tables
        self._max_kv_len = None

        kv_len_host, kv_len_ptr = (None, None)
        if seq_lens is not None:
            kv_len_host, kv_len_ptr = _host_i32_ptr(seq_lens)
            if len(kv_len_host) != batch_size:
                raise ValueError("seq_lens must have batch_size entries")

        # the page table, when it is on the host: the planner inlines every work item's first page
        # ids into its head and (outside CUDA-graph mode) ships the table with the work list
        indices_host = None
        if indices.device.type == "cpu":
            indices_host = indices if indices.is_contiguous() else indices.contiguous()
        self._run_plan(
            indptr_ptr, lpl_ptr, kv_len_ptr, batch_size, num_qo_heads, num_kv_heads, head_dim,
            page_size, q_data_type, kv_data_type, window_left,
            fixed_split_size=fixed_split_size, disable_split_kv=disable_split_kv,
            indices_host=indices_host, head_pairs=self._want_head_pairs(pos_encoding_mode),
            # FI_PLAN_WIDE_UNITS: 16 columns per unit (GQA groups of 9 .. 16, multi-row units); same promises as above
            wide_units=pos_encoding_mode != "ROPE_LLAMA" and self._head_pairs_ok,
        )
        # kept for the one case run() has to plan again (FI_ERR_REPLAN, see _replan_without_head_pairs)
        self._plan_call = ((indptr, indices, last_page_len, num_qo_heads, num_kv_heads, head_dim, page_size),
                           dict(pos_encoding_mode=pos_encoding_mode, window_left=window_left,
                                logits_soft_cap=logits_soft_cap, q_data_type=q_data_type,
                                kv_data_type=kv_data_type, sm_scale=sm_scale, rope_scale=rope_scale,
                                rope_theta=rope_theta, non_blocking=non_blocking, block_tables=block_tables,
                                seq_lens=seq_lens, fixed_split_size=fixed_split_size,
                                disable_split_kv=disable_split_kv))
        if indices_host is not None and not self._use_cuda_graph \
                and self._plan_info[_PI_OFF_INDICES] < 0:
            # the table did not fit behind the work list in the int workspace: our own device copy
            self._indices_dev = indices_host.to(self.device, non_blocking=non_blocking)
        del indptr_host, last_page_len_host, kv_len_host, indices_host  # only needed during the call

        self._pos_encoding_mode = pos_encoding_mode
        self._window_left = window_left
        self._logits_soft_cap = logits_soft_cap
        self._sm_scale = sm_scale
        self._rope_scale = rope_scale
        self._rope_theta = rope_theta
        self._ran_once = False  # step(): the next run() refreshes the run argument block
        # what run(q_len_per_req > 1) needs to plan the per-row requests on first use
        self._mtp_src = (indptr, indices, last_page_len, head_dim, page_size, q_data_type, kv_data_type,
                         fixed_split_size, disable_split_kv)
        if self._use_cuda_graph and self._mtp_children:
            # captured graphs hold the children's kernels: their work lists follow every re-plan
            for n, child in self._mtp_children.items():
                self._plan_mtp_child(child, n)
        else:
            self._mtp_children = {}

    begin_forward = plan

    def _run_multi_token(self, q, paged_kv_cache, q_len_per_req, **run_kw):
        r"""``q_len_per_req`` > 1 (speculative / multi-token decode, reference decode.py:1177,1278;
        SURVEY.md §8(f) rank 4): ``q`` holds ``q_len_per_req`` consecutive query tokens per request
        and the cache already holds their K / V; token ``t`` of a request attends the request's
        first ``kv_len - (q_len_per_req - 1 - t)`` tokens (causal among the new tokens).  A query
        row that attends a prefix of a page list is a decode request over that prefix, so the
        rows run as ``batch_size * q_len_per_req`` planned decode requests on the same kernels.
        The plan of the expanded table is made on the first call (host work) and kept until the
        next ``plan()``.  In CUDA-graph mode the expanded table lives in a child wrapper with its
        own fixed-size buffers that every later ``plan()`` re-plans, so: call
        ``run(q_len_per_req=n)`` once outside the capture (the usual warm-up), capture it, and keep
        re-planning the parent between replays."""
        if self._mtp_src is None:
            raise RuntimeError("q_len_per_req > 1 needs a regular plan() (not fast_decode_plan)")
        if q.shape[0] != self._batch_size * q_len_per_req:
            raise ValueError(
                f"q must hold batch_size * q_len_per_req = {self._batch_size * q_len_per_req} rows, "
                f"got {q.shape[0]}"
            )
        child = self._mtp_children.get(q_len_per_req)
        if child is None:
            if self._use_cuda_graph and torch.cuda.is_current_stream_capturing():
                raise RuntimeError(
                    f"run(q_len_per_req={q_len_per_req}) must be called once before the CUDA-graph capture: "
                    "its per-row plan is host work")
            child = self._make_mtp_child(q_len_per_req)
            self._plan_mtp_child(child, q_len_per_req)
            self._mtp_children[q_len_per_req] = child
        return child.run(q, paged_kv_cache, **run_kw)

    def _mtp_rows_share_the_stream(self) -> bool:
        """Multi-token decode as multi-row units (FiDecodeRunArgs.q_rows): the n rows of a request ride
        its KV stream as extra columns of the GQA group instead of re-reading it once per row — 16-bit
        KV without ROPE_LLAMA / sliding window (the rest takes one decode request per row)."""
        head_dim, page_size, kv_dt = self._mtp_src[3], self._mtp_src[4], self._mtp_src[6]
        return (rows_can_share_the_stream(kv_dt, head_dim, page_size) and self._pos_encoding_mode != "ROPE_LLAMA"
                and self._window_left < 0)

    def _make_mtp_child(self, n: int) -> "BatchDecodeWithPagedKVCacheWrapper":
        if self._mtp_rows_share_the_stream():
            if not self._use_cuda_graph:
                child = BatchDecodeWithPagedKVCacheWrapper(
                    self._float_workspace_buffer, self._kv_layout, use_tensor_cores=self.use_tensor_cores)
            else:  # the request's own table: the parent's capacities
                bs, dev = self._fixed_batch_size, self.device
                child = BatchDecodeWithPagedKVCacheWrapper(
                    self._float_workspace_buffer, self._kv_layout, use_cuda_graph=True,
                    use_tensor_cores=self.use_tensor_cores,
                    paged_kv_indptr_buffer=torch.zeros(bs + 1, dtype=torch.int32, device=dev),
                    paged_kv_indices_buffer=torch.zeros(max(len(self._indices_dev), 1), dtype=torch.int32, device=dev),
                    paged_kv_last_page_len_buffer=torch.zeros(bs, dtype=torch.int32, device=dev))
            child._q_rows = n
            return child
        if not self._use_cuda_graph:
            return BatchDecodeWithPagedKVCacheWrapper(
                self._float_workspace_buffer, self._kv_layout, use_tensor_cores=self.use_tensor_cores)
        # fixed-size buffers for the n-fold expanded table: row t of a request lists a prefix of
        # the request's pages, so n times the parent's capacities always suffice
        bs = self._fixed_batch_size
        dev = self.device
        return BatchDecodeWithPagedKVCacheWrapper(
            self._float_workspace_buffer, self._kv_layout, use_cuda_graph=True,
            use_tensor_cores=self.use_tensor_cores,
            paged_kv_indptr_buffer=torch.zeros(bs * n + 1, dtype=torch.int32, device=dev),
            paged_kv_indices_buffer=torch.zeros(max(len(self._indices_dev) * n, 1), dtype=torch.int32, device=dev),
            paged_kv_last_page_len_buffer=torch.zeros(bs * n, dtype=torch.int32, device=dev))

    def _plan_mtp_child(self, child, n: int) -> None:
        from .cascade import rows_as_decode_requests

        (indptr, indices, last_page_len, head_dim, page_size, q_dt, kv_dt, fss, dskv) = self._mtp_src
        if child._q_rows > 1:  # n x the query heads on the request's own page table
            child.plan(indptr, indices, last_page_len, n * self._num_qo_heads, self._num_kv_heads, head_dim,
                       page_size, pos_encoding_mode=self._pos_encoding_mode, window_left=self._window_left,
                       logits_soft_cap=self._logits_soft_cap, q_data_type=q_dt, kv_data_type=kv_dt,
                       sm_scale=self._sm_scale, rope_scale=self._rope_scale, rope_theta=self._rope_theta,
                       fixed_split_size=fss, disable_split_kv=dskv)
            return
        qo_indptr = torch.arange(self._batch_size + 1, dtype=torch.int32) * n
        r_indptr, r_indices, r_last = rows_as_decode_requests(
            qo_indptr, indptr, indices, last_page_len, page_size, causal=True)
        child.plan(r_indptr, r_indices, r_last, self._num_qo_heads, self._num_kv_heads, head_dim,
                   page_size, pos_encoding_mode=self._pos_encoding_mode, window_left=self._window_left,
                   logits_soft_cap=self._logits_soft_cap, q_data_type=q_dt, kv_data_type=kv_dt,
                   sm_scale=self._sm_scale, rope_scale=self._rope_scale, rope_theta=self._rope_theta,
                   fixed_split_size=fss, disable_split_kv=dskv)

    def _want_head_pairs(self, pos_encoding_mode: str) -> bool:
        """FI_PLAN_HEAD_PAIRS (include/fi_b200.h): an NHD cache keeps the rows of consecutive KV heads
        adjacent, which the fp8 tcgen05 kernel streams as pairs; the planner ignores the bit for every
        shape it does not apply to."""
        if pos_encoding_mode == "ROPE_LLAMA" or not self._head_pairs_ok or self._q_rows > 1:
            return False
        # FI_B200_HEAD_PAIRS=2: also for HND caches (A/B runs; the kernel takes any head stride)
        return self._kv_layout == "NHD" or os.environ.get("FI_B200_HEAD_PAIRS") == "2"

    def _replan_without_head_pairs(self) -> bool:
        """run() found that the head-pair plan does not fit the cache it was given (strides the TMA
        loader cannot address) or the mode (forward() switched to ROPE_LLAMA): plan once more without
        the flag, from the tensors the last plan() was given, and keep it off for this wrapper."""
        if self._plan_call is None or torch.cuda.is_current_stream_capturing():
            return False
        self._head_pairs_ok = False
        args, kw = self._plan_call
        kw = dict(kw, pos_encoding_mode=self._pos_encoding_mode)
        self.plan(*args, **kw)
        return True

    def _bind_workspaces(self) -> None:
        fws, iws, pws = (self._float_workspace_buffer, self._int_workspace_buffer,
                         self._pin_memory_int_workspace_buffer)
        pa, ra = self._pa, self._ra
        pa.float_ws = ra.float_ws = fws.data_ptr()
        pa.float_ws_bytes = fws.numel() * fws.element_size()
        pa.int_ws = ra.int_ws = iws.data_ptr()
        pa.pinned_int_ws = pws.data_ptr()
        pa.int_ws_bytes = min(iws.numel() * iws.element_size(), pws.numel() * pws.element_size())
        pa.device = ra.device = self._dev_idx
        self._bound_ws = (fws, iws, pws)

    def _run_plan(self, indptr_ptr, lpl_ptr, kv_len_ptr, batch_size, num_qo_heads, num_kv_heads,
                  head_dim, page_size, q_data_type, kv_data_type, window_left,
                  fixed_split_size=None, disable_split_kv=False, indices_host=None,
                  head_pairs=False, wide_units=False) -> None:
        # fixed_split_size (pages) / disable_split_kv: batch-invariant planning, reference
        # decode.py:880-887; None / -1 mean "off" there as here
        lib = _lib.load()
        b = self._bound_ws  # engines swap the workspace attributes (reset_workspace_buffer, pokes)
        if b[0] is not self._float_workspace_buffer or b[1] is not self._int_workspace_buffer \
                or b[2] is not self._pin_memory_int_workspace_buffer:
            self._bind_workspaces()
        # a fresh vector per plan (the C function selects the device itself, like the reference's
        # plan, batch_decode.cu:56)
        info = _PlanInfo()
        pa = self._pa
        pa.indptr_host, pa.last_page_len_host, pa.kv_len_host = indptr_ptr, lpl_ptr, kv_len_ptr
        pa.batch_size = batch_size
        # the problem description changes rarely between steps: rewritten only when it does
        sig = (num_qo_heads, num_kv_heads, head_dim, page_size, q_data_type, kv_data_type, window_left,
               fixed_split_size, disable_split_kv, self._use_cuda_graph, head_pairs, wide_units)
        if sig != self._plan_sig:
            pa.flags = (_lib.PLAN_HEAD_PAIRS if head_pairs else 0) | (_lib.PLAN_WIDE_UNITS if wide_units else 0)
            pa.num_qo_heads, pa.num_kv_heads = num_qo_heads, num_kv_heads
            pa.head_dim, pa.page_size = head_dim, page_size
            pa.q_dtype, pa.kv_dtype = dtype_code(q_data_type), dtype_code(kv_data_type)
            pa.window_left = int(window_left)
            pa.cuda_graph = int(self._use_cuda_graph)
            pa.fixed_split_pages = int(fixed_split_size) if fixed_split_size is not None and fixed_split_size > 0 else 0
            pa.disable_split_kv = int(bool(disable_split_kv))
            pa.num_ctas_override = 0
            self._plan_sig = sig
        pa.stream = current_stream_ptr(self.device)
        if indices_host is not None:
            pa.kv_indices_host, pa.num_indices = indices_host.data_ptr(), indices_host.numel()
        else:
            pa.kv_indices_host, pa.num_indices = None, 0
        pa.plan_info_out = C.addressof(info)
        rc = lib.fi_batch_decode_plan_args(self._pa_ref)
        if rc:
            _lib.check(rc)
        self._plan_info = info
        sig = (page_size, num_kv_heads, kv_data_type)
        if sig != self._kv_sig:  # what _bind_cache validated the last cache against has changed
            self._kv_sig, self._kv_key = sig, None
        self._ra.plan_info = C.addressof(info)
        self._head_dim = head_dim
        self._page_size = page_size
        self._q_shape = (batch_size, num_qo_heads, head_dim)
        self._default_sm_scale = 1.0 / math.sqrt(head_dim)
        self._planned_window_left = int(window_left)

    # ------------------------------------------------------------------- run
    def forward(
        self,
        q: torch.Tensor,
        paged_kv_cache: Union[torch.Tensor, Tuple[torch.Tensor, torch.Tensor]],
        pos_encoding_mode: str = "NONE",
        q_scale: Optional[float] = None,
        k_scale: Optional[float] = None,
        v_scale: Optional[float] = None,
        window_left: int = -1,
        logits_soft_cap: Optional[float] = None,
        sm_scale: Optional[float] = None,
        rope_scale: Optional[float] = None,
        rope_theta: Optional[float] = None,
    ) -> torch.Tensor:
        r"""Deprecated alias of :meth:`run` (reference decode.py:1106-1133)."""
        _unsupported_variant(pos_encoding_mode)
        self._pos_encoding_mode = pos_encoding_mode
        self._window_left = window_left
        self._logits_soft_cap = logits_soft_cap
        self._sm_scale = sm_scale
        self._rope_scale = rope_scale
        self._rope_theta = rope_theta
        return self.run(q, paged_kv_cache, q_scale=q_scale, k_scale=k_scale, v_scale=v_scale)

    def run(
        self,
        q: torch.Tensor,
        paged_kv_cache: Union[torch.Tensor, Tuple[torch.Tensor, torch.Tensor]],
        *args,
        q_scale: Optional[float] = None,
        k_scale: Optional[float] = None,
        v_scale: Optional[float] = None,
        out: Optional[torch.Tensor] = None,
        lse: Optional[torch.Tensor] = None,
        return_lse: bool = False,
        enable_pdl: Optional[bool] = None,
        window_left: Optional[int] = None,
        sinks: Optional[torch.Tensor] = None,
        q_len_per_req: Optional[int] = 1,
    ) -> Union[torch.Tensor, Tuple[torch.Tensor, torch.Tensor]]:
        r"""Batch decode attention (reference decode.py:1163-1374).

        q ``[batch_size, num_qo_heads, head_dim]`` (may be a strided view, unit stride on
        head_dim); paged_kv_cache a 5-D tensor ``[pages, 2, page_size, Hkv, D]`` (NHD) /
        ``[pages, 2, Hkv, page_size, D]`` (HND) or a ``(k_cache, v_cache)`` tuple of 4-D
        tensors.  Returns o (q's dtype and shape) and, if ``return_lse``, the base-2
        log-sum-exp ``[batch_size, num_qo_heads]`` (float32).
        """
        if self._plan_info is None:
            raise RuntimeError("plan() must be called before run()")
        if sinks is not None:
            # one extra logit per query head in the softmax denominator (reference decode.py:1176)
            if sinks.dim() != 1 or sinks.shape[0] != self._num_qo_heads // self._q_rows or sinks.device != q.device:
                raise ValueError("sinks must be a [num_qo_heads] tensor on q's device")
            sinks = sinks.to(torch.float32).contiguous()
        if q_len_per_req not in (None, 1):
            return self._run_multi_token(
                q, paged_kv_cache, int(q_len_per_req), q_scale=q_scale, k_scale=k_scale, v_scale=v_scale,
                out=out, lse=lse, return_lse=return_lse, enable_pdl=enable_pdl, window_left=window_left,
                sinks=sinks)
        ra = self._ra
        # ---- the cache: geometry validated once per cache tensor (a serving loop passes the same
        # per-layer tensors every step); tuples and re-allocated caches take the full path
        kv = self._kv_key
        if kv is None or kv[0]() is not paged_kv_cache or kv[1] != paged_kv_cache.data_ptr():
            self._bind_cache(paged_kv_cache)
        # ---- q / out / lse
        if q.dtype != self._cached_q_data_type:
            raise ValueError(
                f"The dtype of q {q.dtype} does not match the q_data_type {self._cached_q_data_type} "
                "specified in plan function.")
        qs = q.stride()
        if q.shape != self._q_shape:
            rows = self._q_rows  # a multi-token child: [batch * rows, Hq, D] against a plan of rows * Hq heads
            if rows == 1 or q.shape != (self._batch_size * rows, self._num_qo_heads // rows, self._head_dim):
                raise ValueError(
                    f"q must have shape [{self._batch_size * rows}, {self._num_qo_heads // rows}, {self._head_dim}], "
                    f"got {list(q.shape)}"
                )
        if qs[2] != 1:
            raise ValueError("the last dimension of q / k / v must be contiguous")
        dev = q.device
        if dev != self.device:
            raise ValueError("q / paged_kv_cache must be on the wrapper's device")

        # window_left is baked into the plan (the segments' token bounds): a different value at run
        # time cannot be honoured (the reference passes it to its kernel, decode.py:1295)
        if (self._window_left if window_left is None else window_left) != self._planned_window_left:
            raise ValueError(
                f"window_left {self._window_left if window_left is None else window_left} differs from "
                f"the planned one ({self._planned_window_left}): call plan() again")
        logits_soft_cap = self._logits_soft_cap
        sm_scale = self._sm_scale
        if sm_scale is None:
            sm_scale = self._default_sm_scale
        if q_scale is not None:
            sm_scale *= q_scale
        if k_scale is not None:
            sm_scale *= k_scale

        if return_lse:
            if lse is None:
                lse = torch.empty((q.shape[0], q.shape[1]), dtype=torch.float32, device=dev)
            else:
                check_shape_dtype_device(lse, (q.size(0), q.size(1)), torch.float32, dev, "lse")
                if not lse.is_contiguous():
                    raise ValueError("lse must be contiguous")
            ra.lse = lse.data_ptr()
        else:
            ra.lse = None
        if out is None:
            out = torch.empty(q.shape, dtype=q.dtype, device=dev)
        elif out.shape != q.shape or out.dtype != q.dtype or out.device != dev:
            # beyond the reference: `out` may be a PINNED host tensor — the kernel's epilogue stores o
            # straight into it over PCIe (pinned memory is device-addressable under UVA), so the
            # device-to-host copy of the result overlaps the kernel instead of following it.  The
            # caller synchronizes the stream before reading it, as after any non_blocking copy.
            if not (out.device.type == "cpu" and out.shape == q.shape and out.dtype == q.dtype
                    and out.is_contiguous() and out.is_pinned()):
                check_shape_dtype_device(out, q.shape, q.dtype, dev, "out")
                raise ValueError("a host `out` must be pinned and contiguous")
        elif not out.is_contiguous():
            raise ValueError("out must be contiguous")

        pk = (self._pos_encoding_mode, self._rope_scale, self._rope_theta)
        if pk != self._pos_key:
            (ra.pos_encoding_mode, ra.alibi_slopes, ra.rope_scale, ra.rope_theta) = _pos_mode_args(
                self._pos_encoding_mode, q.shape[1], dev, self._rope_scale, self._rope_theta)
            self._pos_key = pk
        ra.q, ra.q_stride_n, ra.q_stride_h = q.data_ptr(), qs[0], qs[1]
        ra.kv_indices = self._indices_dev.data_ptr() if self._indices_dev is not None else None
        ra.kv_indptr = self._indptr_dev.data_ptr() if self._indptr_dev is not None else None
        ra.o = out.data_ptr()
        ra.sm_scale = sm_scale
        # v_scale is fused into the kernel epilogue (the reference multiplies afterwards,
        # decode.py:1367-1372)
        ra.v_scale = v_scale if v_scale is not None else 1.0
        ra.logits_soft_cap = logits_soft_cap if logits_soft_cap is not None else 0.0
        ra.sinks = sinks.data_ptr() if sinks is not None else None
        ra.enable_pdl = self._pdl_default if enable_pdl is None else int(bool(enable_pdl))
        ra.q_rows = self._q_rows
        ra.q_rows_noncausal = self._q_rows_noncausal
        ra.stream = current_stream_ptr(dev)
        rc = self._lib.fi_batch_decode_run_args(self._ra_ref)
        if rc:
            if rc == _lib.FI_ERR_REPLAN and self._replan_without_head_pairs():
                rc = self._lib.fi_batch_decode_run_args(self._ra_ref)
            if rc:
                _lib.check(rc)
        self._ran_once = True
        self._run_kw = (q_scale, k_scale, v_scale, enable_pdl, sinks)
        return (out, lse) if return_lse else out

    def step(self, indptr: torch.Tensor, indices: torch.Tensor, last_page_len: torch.Tensor,
             q: torch.Tensor, paged_kv_cache, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        r"""One decode step of a serving loop in ONE library call (beyond the reference, whose wrapper
        plans and runs through separate calls): ``plan(indptr, indices, last_page_len, <everything
        else as in the last plan()>)`` followed by ``run(q, paged_kv_cache, out=out)`` with the
        scales, soft cap, sinks and PDL setting of the last ``run()``.

        The fast path needs what a serving loop has anyway: a previous ``plan()`` and ``run()`` on this
        wrapper, the same cache tensor as that ``run()``, and the page table as contiguous int32 CPU
        tensors (pinned or not; the planner reads them during the call).  Anything else — CUDA-graph
        mode, device tables, another cache — takes ``plan()`` + ``run()``.
        """
        kv = self._kv_key
        call = self._plan_call
        i32 = torch.int32
        if (call is None or self._use_cuda_graph or kv is None or not self._ran_once
                or kv[0]() is not paged_kv_cache or kv[1] != paged_kv_cache.data_ptr()
                or indptr.device.type != "cpu" or indices.device.type != "cpu" or last_page_len.device.type != "cpu"
                or indptr.dtype != i32 or indices.dtype != i32 or last_page_len.dtype != i32
                or not (indptr.is_contiguous() and indices.is_contiguous() and last_page_len.is_contiguous())):
            if call is None:
                raise RuntimeError("step() needs a previous plan() on this wrapper")
            self.plan(indptr, indices, last_page_len, *call[0][3:], **dict(call[1], seq_lens=None, block_tables=None))
            rk = self._run_kw  # the last run()'s scales, as on the fast path
            return self.run(q, paged_kv_cache, out=out, q_scale=rk[0], k_scale=rk[1], v_scale=rk[2],
                            enable_pdl=rk[3], sinks=rk[4])
        batch_size = last_page_len.shape[0]
        if indptr.shape[0] != batch_size + 1:
            raise ValueError("indptr must have batch_size + 1 entries")
        if q.dtype != self._cached_q_data_type:
            raise ValueError(
                f"The dtype of q {q.dtype} does not match the q_data_type {self._cached_q_data_type} "
                "specified in plan function.")
        shape = (batch_size, self._num_qo_heads, self._head_dim)
        qs = q.stride()
        if q.shape != shape or qs[2] != 1 or q.device != self.device:
            raise ValueError(f"q must be a [{batch_size}, {self._num_qo_heads}, {self._head_dim}] tensor on the "
                             "wrapper's device with a contiguous last dimension")
        if out is None:
            out = torch.empty(shape, dtype=q.dtype, device=q.device)
        elif out.shape != q.shape or out.dtype != q.dtype or not out.is_contiguous() or \
                not (out.device == q.device or (out.device.type == "cpu" and out.is_pinned())):
            raise ValueError("out must be a contiguous tensor of q's shape and dtype on q's device or in pinned host memory")
        pa, ra = self._pa, self._ra
        info = _PlanInfo()
        pa.indptr_host, pa.last_page_len_host, pa.kv_len_host = indptr.data_ptr(), last_page_len.data_ptr(), None
        pa.kv_indices_host, pa.num_indices = indices.data_ptr(), indices.numel()
        pa.batch_size = batch_size
        pa.plan_info_out = ra.plan_info = C.addressof(info)
        pa.stream = ra.stream = current_stream_ptr(self.device)
        ra.q, ra.q_stride_n, ra.q_stride_h = q.data_ptr(), qs[0], qs[1]
        ra.o, ra.lse = out.data_ptr(), None
        ra.kv_indices = ra.kv_indptr = None  # the planner ships the table behind the work list
        rc = self._lib.fi_batch_decode_step(self._pa_ref, self._ra_ref)
        self._plan_info = info
        if rc:
            if rc in (_lib.FI_ERR_REPLAN, _lib.FI_ERR_WORKSPACE):  # head pairs / table too large: the long way
                self._ran_once = False
                try:
                    return self.step(indptr, indices, last_page_len, q, paged_kv_cache, out=out)
                finally:
                    self._ran_once = True
            _lib.check(rc)
        # what plan() leaves behind for the other entry points
        self._indices_dev, self._indices_src = None, indices
        self._indptr_dev, self._indptr_src = None, indptr
        self._last_page_len_dev, self._last_page_len_src = None, last_page_len
        self._batch_size, self._q_shape = batch_size, shape
        self._plan_call = ((indptr, indices, last_page_len) + call[0][3:], call[1])
        self._mtp_src = (indptr, indices, last_page_len) + self._mtp_src[3:]
        if self._mtp_children:
            self._mtp_children = {}
        return out

    def _bind_cache(self, paged_kv_cache) -> None:
        """Validate the cache against the plan and write its geometry into the run argument block."""
        import weakref

        if torch.is_tensor(paged_kv_cache) and paged_kv_cache.dim() == 5:
            # the common 5-D cache: K and V are the two halves of dim 1 (flashinfer/utils.py:160-162
            # unbinds them) — same strides by construction, no view objects on the hot path
            c = paged_kv_cache
            cs = c.stride()
            kv_dtype = c.dtype
            kv_dev = c.device
            k_ptr = c.data_ptr()
            v_ptr = k_ptr + cs[1] * c.element_size()
            num_pages, stride_page, last_stride = c.shape[0], cs[0], cs[4]
            d2, d3, s2, s3 = c.shape[2], c.shape[3], cs[2], cs[3]
            key = (weakref.ref(c), k_ptr)
        else:
            k_cache, v_cache = _unpack_paged_kv_cache(paged_kv_cache, self._kv_layout)
            if v_cache.dtype != k_cache.dtype:
                raise ValueError("k_cache and v_cache must have the same dtype")
            if k_cache.stride() != v_cache.stride():
                raise ValueError("k/v strides must be identical")
            ks = k_cache.stride()
            kv_dtype, kv_dev = k_cache.dtype, k_cache.device
            k_ptr, v_ptr = k_cache.data_ptr(), v_cache.data_ptr()
            num_pages, stride_page, last_stride = k_cache.shape[0], ks[0], ks[3]
            d2, d3, s2, s3 = k_cache.shape[1], k_cache.shape[2], ks[1], ks[2]
            key = None  # tuples / views are rebuilt by the caller on every call: nothing to key on
        self._kv_key = None
        if kv_dtype != self._cached_kv_data_type:
            raise ValueError(
                f"The dtype of k {kv_dtype} does not match the kv_data_type {self._cached_kv_data_type} "
                "specified in plan function.")
        if self._kv_layout == "NHD":
            page_size, num_kv_heads, stride_n, stride_h = d2, d3, s2, s3
        else:
            num_kv_heads, page_size, stride_h, stride_n = d2, d3, s2, s3
        if page_size != self._page_size or num_kv_heads != self._num_kv_heads:
            raise ValueError(
                "paged_kv_cache shape does not match the (page_size, num_kv_heads) given to plan()"
            )
        if last_stride != 1:
            raise ValueError("the last dimension of q / k / v must be contiguous")
        if kv_dev != self.device:
            raise ValueError("q / paged_kv_cache must be on the wrapper's device")
        ra = self._ra
        ra.k_cache, ra.v_cache, ra.num_pages = k_ptr, v_ptr, num_pages
        ra.kv_stride_page, ra.kv_stride_n, ra.kv_stride_h = stride_page, stride_n, stride_h
        self._kv_key = key

    def forward_return_lse(
        self,
        q: torch.Tensor,
        paged_kv_cache: torch.Tensor,
        pos_encoding_mode: str = "NONE",
        q_scale: Optional[float] = None,
        k_scale: Optional[float] = None,
        v_scale: Optional[float] = None,
        window_left: int = -1,
        logits_soft_cap: Optional[float] = None,
        sm_scale: Optional[float] = None,
        rope_scale: Optional[float] = None,
        rope_theta: Optional[float] = None,
    ) -> Tuple[torch.Tensor, torch.Tensor]:
        r"""Deprecated alias of :meth:`run_return_lse` (reference decode.py:1376-1403)."""
        _unsupported_variant(pos_encoding_mode)
        self._pos_encoding_mode = pos_encoding_mode
        self._window_left = window_left
        self._logits_soft_cap = logits_soft_cap
        self._sm_scale = sm_scale
        self._rope_scale = rope_scale
        self._rope_theta = rope_theta
        return self.run(q, paged_kv_cache, q_scale=q_scale, k_scale=k_scale, v_scale=v_scale,
                        return_lse=True)

    run_return_lse = functools.partialmethod(run, return_lse=True)

    def end_forward(self) -> None:
        r"""Deprecated, no effect (reference decode.py:1407-1409)."""
        pass


class CUDAGraphBatchDecodeWithPagedKVCacheWrapper(BatchDecodeWithPagedKVCacheWrapper):
    r"""CUDA-graph flavour with user-owned index buffers (reference decode.py:1413-1480)."""

    def __init__(
        self,
        workspace_buffer: torch.Tensor,
        indptr_buffer: torch.Tensor,
        indices_buffer: torch.Tensor,
        last_page_len_buffer: torch.Tensor,
        kv_layout: str = "NHD",
        use_tensor_cores: bool = False,
    ) -> None:
        super().__init__(
            workspace_buffer,
            kv_layout,
            use_cuda_graph=True,
            use_tensor_cores=use_tensor_cores,
            paged_kv_indptr_buffer=indptr_buffer,
            paged_kv_indices_buffer=indices_buffer,
            paged_kv_last_page_len_buffer=last_page_len_buffer,
        )


def fast_decode_plan(
    self: BatchDecodeWithPagedKVCacheWrapper,
    indptr: torch.Tensor,
    indices: torch.Tensor,
    last_page_len: torch.Tensor,
    num_qo_heads: int,
    num_kv_heads: int,
    head_dim: int,
    page_size: int,
    pos_encoding_mode: str = "NONE",
    window_left: int = -1,
    logits_soft_cap: Optional[float] = None,
    q_data_type: Optional[Union[str, torch.dtype]] = None,
    kv_data_type: Optional[Union[str, torch.dtype]] = None,
    data_type: Optional[Union[str, torch.dtype]] = None,
    sm_scale: Optional[float] = None,
    rope_scale: Optional[float] = None,
    rope_theta: Optional[float] = None,
    non_blocking: bool = True,
    fixed_split_size: Optional[int] = None,
    disable_split_kv: bool = False,
    global_override_indptr_cpu: Optional[torch.Tensor] = None,
    last_page_len_cpu: Optional[torch.Tensor] = None,
) -> None:
    r"""Lighter ``plan`` that adopts the caller's device index tensors without copying
    (reference decode.py:2416-2579).  ``global_override_indptr_cpu`` avoids the D2H read of
    ``indptr``; ``last_page_len`` is read on the host unless ``page_size == 1`` or the caller hands
    over its host copy as ``last_page_len_cpu`` (an addition to the reference's signature: the token-
    exact planner needs the lengths, and a serving engine has them on the host anyway), which makes
    the call free of device synchronisation.
    """
    batch_size = last_page_len.shape[0]
    if logits_soft_cap is None:
        logits_soft_cap = 0.0
    _unsupported_variant(pos_encoding_mode)
    if data_type is not None:
        if q_data_type is None:
            q_data_type = data_type
        if kv_data_type is None:
            kv_data_type = data_type
    if q_data_type is None:
        q_data_type = getattr(self, "_cached_q_data_type", None) or torch.float16
    q_data_type = canonicalize_torch_dtype(q_data_type)
    if kv_data_type is None:
        kv_data_type = q_data_type
    kv_data_type = canonicalize_torch_dtype(kv_data_type)

    if self.is_cuda_graph_enabled:
        if batch_size != self._fixed_batch_size:
            raise ValueError(
                "The batch size should be fixed in cudagraph mode, the runtime batch size {} "
                " mismatches the batch size set during initialization {}".format(
                    batch_size, self._fixed_batch_size
                )
            )
        if len(indices) > len(self._paged_kv_indices_buf):
            raise ValueError(
                "The size of indices should be less than or equal to the allocated buffer"
            )
    else:
        self._paged_kv_indptr_buf = indptr
        self._paged_kv_indices_buf = indices
        self._paged_kv_last_page_len_buf = last_page_len

    indptr_src = global_override_indptr_cpu if global_override_indptr_cpu is not None else indptr
    indptr_host, indptr_ptr = _host_i32_ptr(indptr_src)
    if last_page_len_cpu is not None:
        last_page_len_host, _ = _host_i32_ptr(last_page_len_cpu)
        if last_page_len_host.shape[0] != batch_size:
            raise ValueError("last_page_len_cpu must have batch_size entries")
    elif page_size == 1:
        last_page_len_host = torch.ones((batch_size,), dtype=torch.int32)
    else:
        last_page_len_host = last_page_len.detach().to("cpu", torch.int32).contiguous()
    self._cached_q_data_type = q_data_type
    self._cached_kv_data_type = kv_data_type
    self._batch_size = batch_size
    self._num_qo_heads = num_qo_heads
    self._num_kv_heads = num_kv_heads
    try:
        self._plan_call = None
        self._run_plan(indptr_ptr, last_page_len_host.data_ptr(), None, batch_size, num_qo_heads,
                       num_kv_heads, head_dim, page_size, q_data_type, kv_data_type, window_left,
                       fixed_split_size=fixed_split_size, disable_split_kv=disable_split_kv,
                       head_pairs=self._want_head_pairs(pos_encoding_mode),
                       wide_units=pos_encoding_mode != "ROPE_LLAMA" and self._head_pairs_ok)
    except Exception as e:
        raise RuntimeError(f"Error in standard plan: {e}") from e
    self._pos_encoding_mode = pos_encoding_mode
    self._window_left = window_left
    self._logits_soft_cap = logits_soft_cap
    self._sm_scale = sm_scale
    self._rope_scale = rope_scale
    self._rope_theta = rope_theta
    self._mtp_children = {}
    self._mtp_src = (indptr_src, indices, last_page_len, head_dim, page_size, q_data_type, kv_data_type,
                     fixed_split_size, disable_split_kv)
