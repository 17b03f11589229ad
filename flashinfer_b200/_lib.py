"""ctypes binding of libfi_b200.so — the thin C-ABI boundary (include/fi_b200.h).

The reference loads one tvm-ffi module per (dtype, head_dim, variant) through its JIT
(flashinfer/jit/core.py:247-263, flashinfer/decode.py:208-302).  Here there is ONE
prebuilt shared library with plain-C entry points; this module declares their
prototypes and converts non-zero status codes into the Python exceptions the
reference raises (ValueError for argument errors, RuntimeError otherwise).

There is deliberately no fallback: if the library is missing, importing the ops
raises, it never reroutes to a CPU / PyTorch path.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

_HERE = Path(__file__).resolve().parent
LIB_PATH = _HERE / "libfi_b200.so"

FI_OK = 0
FI_ERR_INVALID_ARGUMENT = 1
FI_ERR_WORKSPACE = 2
FI_ERR_CUDA = 3
FI_ERR_UNSUPPORTED = 4
FI_ERR_REPLAN = 5
PLAN_HEAD_PAIRS = 1
PLAN_WIDE_UNITS = 2

DTYPE_F16, DTYPE_BF16, DTYPE_E4M3, DTYPE_E5M2, DTYPE_F32 = 0, 1, 2, 3, 4
LAYOUT_NHD, LAYOUT_HND = 0, 1

_vp, _i, _i64, _f, _sz = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_size_t

_i32, _u32 = C.c_int32, C.c_uint32


class DecodePlanArgs(C.Structure):
    """FiDecodePlanArgs (include/fi_b200.h), field for field."""
    _fields_ = [("float_ws", _vp), ("float_ws_bytes", _sz), ("int_ws", _vp), ("pinned_int_ws", _vp),
                ("int_ws_bytes", _sz), ("indptr_host", _vp), ("last_page_len_host", _vp),
                ("kv_len_host", _vp), ("batch_size", _i32), ("num_qo_heads", _i32),
                ("num_kv_heads", _i32), ("head_dim", _i32), ("page_size", _i32), ("q_dtype", _i32),
                ("kv_dtype", _i32), ("window_left", _i32), ("cuda_graph", _i32),
                ("fixed_split_pages", _i32), ("disable_split_kv", _i32), ("device", _i32),
                ("num_ctas_override", _i32), ("flags", _i32), ("stream", _vp), ("kv_indices_host", _vp),
                ("num_indices", _i64), ("plan_info_out", _vp)]


class DecodeRunArgs(C.Structure):
    """FiDecodeRunArgs (include/fi_b200.h), field for field."""
    _fields_ = [("float_ws", _vp), ("int_ws", _vp), ("plan_info", _vp), ("q", _vp),
                ("q_stride_n", _i64), ("q_stride_h", _i64), ("k_cache", _vp), ("v_cache", _vp),
                ("num_pages", _i64), ("kv_stride_page", _i64), ("kv_stride_n", _i64),
                ("kv_stride_h", _i64), ("kv_indices", _vp), ("kv_indptr", _vp), ("o", _vp), ("lse", _vp),
                ("sm_scale", _f), ("v_scale", _f), ("logits_soft_cap", _f), ("pos_encoding_mode", _i32),
                ("alibi_slopes", _vp), ("rope_scale", _f), ("rope_theta", _f), ("sinks", _vp),
                ("enable_pdl", _i32), ("device", _i32), ("stream", _vp), ("q_rows", _i32), ("q_rows_noncausal", _i32)]


# name -> (restype, argtypes); must list every symbol include/fi_b200.h declares
PROTOTYPES = {
    "fi_last_error": (C.c_char_p, []),
    "fi_version": (_i, []),
    "fi_plan_info_len": (_i, []),
    "fi_set_tuning": (None, [_i, _i, _i, _i]),
    "fi_last_loader": (_i, []),
    "fi_launch_count": (C.c_longlong, []),
    "fi_batch_decode_plan": (_i, [_vp, _sz, _vp, _vp, _sz, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i,
                                  _i, _i, _i, _i, _i, _i, _i, _vp, _vp, _i64, _vp]),
    "fi_batch_decode_run": (_i, [_vp, _vp, _vp, _vp, _i64, _i64, _vp, _vp, _i64, _i64, _i64, _i64,
                                 _vp, _vp, _vp, _vp, _f, _f, _f, _i, _vp, _f, _f, _vp, _i, _i, _vp]),
    "fi_batch_decode_plan_args": (_i, [C.POINTER(DecodePlanArgs)]),
    "fi_batch_decode_run_args": (_i, [C.POINTER(DecodeRunArgs)]),
    "fi_batch_decode_step": (_i, [C.POINTER(DecodePlanArgs), C.POINTER(DecodeRunArgs)]),
    "fi_single_decode_run": (_i, [_vp, _vp, _vp, _vp, _sz, _vp, _vp, _i, _i, _i, _i, _i, _i, _i,
                                  _i, _f, _f, _f, _i, _vp, _f, _f, _i, _vp]),
    "fi_single_decode_tmp_bytes": (_sz, [_i, _i, _i, _i]),
    "fi_single_decode_run_static": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i,
                                         _i, _f, _f, _f, _i, _vp, _f, _f, _i, _vp]),
    "fi_merge_state": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "fi_merge_state_in_place": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "fi_merge_states": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    "fi_variable_length_merge_states": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _f,
                                             _i, _vp]),
    "fi_get_batch_indices_positions": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _vp]),
    "fi_peer_gather": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i64, _i64, _i64, _i64, C.c_uint32, _i, _i, _vp]),
    "fi_append_paged_kv_cache": (_i, [_vp, _vp, _i64, _i64, _i64, _i64, _vp, _vp, _i64, _vp, _vp,
                                      _i64, _i64, _i64, _vp, _vp, _i, _i, _i, _i, _i, _vp]),
}

_lib = None


def load() -> C.CDLL:
    """Load libfi_b200.so (building it first if FI_B200_AUTOBUILD=1 and it is absent)."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        if os.environ.get("FI_B200_AUTOBUILD", "0") == "1":
            from .build import build

            build()
        else:
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -m flashinfer_b200.build` "
                "(there is no CPU / PyTorch fallback for the decode path)"
            )
    # FI_B200_LIB: load another build of the same ABI (the -DFI_B200_TRACE one, scripts/trace_probe.py)
    lib = C.CDLL(os.environ.get("FI_B200_LIB", str(LIB_PATH)))
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError if the .so does not export it
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def last_error() -> str:
    msg = load().fi_last_error()
    return msg.decode("utf-8", "replace") if msg else ""


def check(status: int) -> None:
    """Map a C status code to the reference's exception classes."""
    if status == FI_OK:
        return
    msg = last_error()
    if status == FI_ERR_INVALID_ARGUMENT:
        raise ValueError(msg)
    raise RuntimeError(msg)
