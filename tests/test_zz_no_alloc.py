"""fi_single_decode_run_static — the allocation-free entry point (include/fi_b200.h).

Collected last on purpose (`test_zz_`): a failure here must
not hide any other test.  Called through the C-ABI (ctypes), compared with the existing entry
point fi_single_decode_run on the same inputs, on two calls in a row (the second call finds the
first call's partials in the static scratch), and with the CPU oracle.
"""
from __future__ import annotations

import pytest
import torch

from oracle import attention_ref as orc

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
O_TOL, LSE_TOL = 1e-2, 1e-3  # the gates every other parity test in this suite uses


def _args(q, k, v, o, lse, hq, hkv, d, kv_len, layout=0):
    from flashinfer_b200.utils import current_stream_ptr, dtype_code

    return (o.data_ptr(), lse.data_ptr(), hq, hkv, d, kv_len, dtype_code(q.dtype), dtype_code(k.dtype),
            layout, -1, float(d) ** -0.5, 1.0, 0.0, 0, None, 1.0, 1e4, 0, current_stream_ptr(torch.device(DEV)))


def _both(lib, q, k, v, hq, hkv, d, kv_len):
    """(o, lse) of the static-scratch entry point and of the existing one, same device inputs."""
    from flashinfer_b200 import _lib

    qd, kd, vd = q.to(DEV), k.to(DEV), v.to(DEV)
    o_s, lse_s = torch.empty_like(qd), torch.empty(hq, dtype=torch.float32, device=DEV)
    o_e, lse_e = torch.empty_like(qd), torch.empty(hq, dtype=torch.float32, device=DEV)
    tmp_bytes = lib.fi_single_decode_tmp_bytes(hq, hkv, d, 0)
    tmp = torch.empty(tmp_bytes, dtype=torch.uint8, device=DEV)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info(0)
    alloc0 = torch.cuda.memory_allocated(0)
    _lib.check(lib.fi_single_decode_run_static(qd.data_ptr(), kd.data_ptr(), vd.data_ptr(),
                                               *_args(qd, kd, vd, o_s, lse_s, hq, hkv, d, kv_len)))
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(0)
    assert torch.cuda.memory_allocated(0) == alloc0
    _lib.check(lib.fi_single_decode_run(qd.data_ptr(), kd.data_ptr(), vd.data_ptr(), tmp.data_ptr(), tmp.numel(),
                                        *_args(qd, kd, vd, o_e, lse_e, hq, hkv, d, kv_len)))
    torch.cuda.synchronize()
    return (o_s, lse_s), (o_e, lse_e), free0 - free1


def test_static_scratch_single_decode_matches_existing_entry_point_twice():
    from flashinfer_b200 import _lib

    lib = _lib.load()
    hq, hkv, d = 32, 8, 128
    grew = []
    # two calls in a row with different lengths and data: the second one runs over the first one's
    # partials (longer first, so stale slots beyond the second call's need stay in the array)
    for call, kv_len in enumerate((8192, 2048, 777)):
        g = torch.Generator().manual_seed(100 + call)
        q = torch.randn(hq, d, generator=g).half()
        k = torch.randn(kv_len, hkv, d, generator=g).half()
        v = torch.randn(kv_len, hkv, d, generator=g).half()
        (o_s, lse_s), (o_e, lse_e), used = _both(lib, q, k, v, hq, hkv, d, kv_len)
        grew.append(used)
        assert (o_s.float() - o_e.float()).abs().max().item() <= O_TOL
        assert (lse_s - lse_e).abs().max().item() <= LSE_TOL
        # same kernels, same split, same merge order: in fact bit-identical
        assert torch.equal(o_s, o_e) and torch.equal(lse_s, lse_e)
        o_ref, lse_ref = orc.decode_attention_ref(q, k, v)
        assert (o_s.float().cpu() - o_ref).abs().max().item() <= O_TOL
        assert (lse_s.cpu() - lse_ref).abs().max().item() <= LSE_TOL
    # once the module is resident (first call), a call takes no device memory at all
    assert grew[1] <= 0 and grew[2] <= 0, grew


def test_static_scratch_rejects_shapes_beyond_its_maximum():
    from flashinfer_b200 import _lib

    lib = _lib.load()
    hq, hkv, d, kv_len = 256, 8, 256, 256  # needs ~10 MB of partial slots
    assert lib.fi_single_decode_tmp_bytes(hq, hkv, d, 0) > (4 << 20)
    q = torch.zeros(hq, d, dtype=torch.float16, device=DEV)
    k = torch.zeros(kv_len, hkv, d, dtype=torch.float16, device=DEV)
    o, lse = torch.empty_like(q), torch.empty(hq, dtype=torch.float32, device=DEV)
    rc = lib.fi_single_decode_run_static(q.data_ptr(), k.data_ptr(), k.data_ptr(),
                                         *_args(q, k, k, o, lse, hq, hkv, d, kv_len))
    assert rc == _lib.FI_ERR_WORKSPACE
    msg = _lib.last_error()
    assert "static" in msg and "FI_SINGLE_DECODE_STATIC_TMP_BYTES" in msg, msg
