"""GPU parity of the attention-state merge operators against the closed-form oracle
(reference tests: tests/utils/test_triton_cascade.py:9-86, atol 1e-2;
tests/attention/test_shared_prefix_kernels.py:229-310 mask semantics)."""
import pytest
import torch

from oracle import attention_ref as orc

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _fi():
    import flashinfer_b200 as fi

    return fi


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16, torch.float32])
@pytest.mark.parametrize("seq_len,heads,d", [(2048, 32, 128), (7, 3, 64), (1, 1, 256)])
def test_merge_state(dtype, seq_len, heads, d):
    g = torch.Generator().manual_seed(0)
    va = torch.randn(seq_len, heads, d, generator=g).to(dtype)
    vb = torch.randn(seq_len, heads, d, generator=g).to(dtype)
    sa = torch.randn(seq_len, heads, generator=g) * 3
    sb = torch.randn(seq_len, heads, generator=g) * 3
    v, s = _fi().merge_state(va.to(DEV), sa.to(DEV), vb.to(DEV), sb.to(DEV))
    v_ref, s_ref = orc.merge_state_ref(va, sa, vb, sb)
    assert v.dtype == dtype and s.dtype == torch.float32
    # atol = rtol = 1e-2 as in tests/utils/test_triton_cascade.py:21-22 (bf16 has 8 mantissa bits)
    torch.testing.assert_close(v.float().cpu(), v_ref, atol=1e-2, rtol=1e-2)
    assert (s.cpu() - s_ref).abs().max() <= 1e-3


@pytest.mark.parametrize("use_mask", [False, True])
def test_merge_state_in_place(use_mask):
    g = torch.Generator().manual_seed(1)
    seq_len, heads, d = 333, 8, 128
    v0 = torch.randn(seq_len, heads, d, generator=g).half()
    vo = torch.randn(seq_len, heads, d, generator=g).half()
    s0 = torch.randn(seq_len, heads, generator=g)
    so = torch.randn(seq_len, heads, generator=g)
    mask = (torch.rand(seq_len, generator=g) > 0.5) if use_mask else None
    v, s = v0.clone().to(DEV), s0.clone().to(DEV)
    _fi().merge_state_in_place(v, s, vo.to(DEV), so.to(DEV), mask=None if mask is None else mask.to(DEV))
    v_ref, s_ref = orc.merge_state_ref(v0, s0, vo, so)
    if mask is not None:
        keep = ~mask
        v_ref[keep] = v0.float()[keep]
        s_ref[keep] = s0[keep]
        assert torch.equal(v.cpu()[keep], v0[keep])  # untouched rows are bit-identical
    torch.testing.assert_close(v.float().cpu(), v_ref, atol=1e-2, rtol=1e-2)
    assert (s.cpu() - s_ref).abs().max() <= 1e-3


@pytest.mark.parametrize("n", [1, 2, 7, 100])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_merge_states(n, dtype):
    g = torch.Generator().manual_seed(n)
    seq_len, heads, d = 129, 32, 128
    v = torch.randn(seq_len, n, heads, d, generator=g).to(dtype)
    s = torch.randn(seq_len, n, heads, generator=g) * 4
    vm, sm = _fi().merge_states(v.to(DEV), s.to(DEV))
    v_ref, s_ref = orc.merge_states_ref(v, s)
    torch.testing.assert_close(vm.float().cpu(), v_ref, atol=1e-2, rtol=1e-2)
    assert (sm.cpu() - s_ref).abs().max() <= 1e-3


def test_variable_length_merge_states():
    g = torch.Generator().manual_seed(9)
    counts = [3, 0, 1, 17, 2]
    indptr = torch.tensor([0] + list(torch.tensor(counts).cumsum(0)), dtype=torch.int32)
    total, heads, d = sum(counts), 8, 128
    v = torch.randn(total, heads, d, generator=g)
    s = torch.randn(total, heads, generator=g) * 3
    vm, sm = _fi().variable_length_merge_states(v.to(DEV), s.to(DEV), indptr, out_dtype=torch.bfloat16)
    v_ref, s_ref = orc.variable_length_merge_states_ref(v, s, indptr.tolist())
    torch.testing.assert_close(vm.float().cpu(), v_ref, atol=1e-2, rtol=1e-2)
    assert (sm.cpu() - s_ref).abs().max() <= 1e-3
    assert torch.all(vm[1] == 0) and torch.all(sm[1] == -5e4)


def test_merge_is_associative_with_decode_partials():
    """merge_state(merge_state(a, b), c) == merge_states([a, b, c])."""
    g = torch.Generator().manual_seed(2)
    vs = torch.randn(64, 3, 4, 128, generator=g).half().to(DEV)
    ss = (torch.randn(64, 3, 4, generator=g) * 2).to(DEV)
    fi = _fi()
    v_ab, s_ab = fi.merge_state(vs[:, 0], ss[:, 0], vs[:, 1], ss[:, 1])
    v1, s1 = fi.merge_state(v_ab, s_ab, vs[:, 2], ss[:, 2])
    v2, s2 = fi.merge_states(vs, ss)
    assert (v1.float() - v2.float()).abs().max() <= 4e-3
    assert (s1 - s2).abs().max() <= 1e-4
