"""Generate tests/golden/alibi_golden.npz by EXECUTING the reference's own ALiBi oracle
(/root/reference/tests/test_helpers/alibi_reference.py: `alibi_attention`, the function its
tests/attention/test_alibi.py:55-70 compares single_decode_with_kv_cache(pos_encoding_mode="ALIBI")
against).  Build container only (reads /root/reference):  python tests/golden/make_golden_alibi.py

The module is pure PyTorch and is loaded by path, unchanged.  Inputs are stored with the outputs.
"""
import importlib.util
from pathlib import Path

import numpy as np
import torch

REF = Path("/root/reference/tests/test_helpers/alibi_reference.py")
OUT = Path(__file__).resolve().parent / "alibi_golden.npz"

# (name, num_heads, head_dim, seq_len, dtype) — the reference test's grid (seq_len 1, 9, 81, 729;
# heads 4, 8, 32; head_dim 128, 256) thinned to what is small enough to commit, plus a
# non-power-of-two head count for the second slope family
CASES = [
    ("h4_d128_len1", 4, 128, 1, torch.float16),
    ("h4_d128_len9", 4, 128, 9, torch.float16),
    ("h8_d128_len81", 8, 128, 81, torch.float16),
    ("h8_d256_len81", 8, 256, 81, torch.float16),
    ("h12_d128_len70_bf16", 12, 128, 70, torch.bfloat16),
    ("h32_d128_len33", 32, 128, 33, torch.float16),
]


def main():
    spec = importlib.util.spec_from_file_location("alibi_reference", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    out = {"names": np.array([c[0] for c in CASES])}
    for i, (name, H, D, L, dtype) in enumerate(CASES):
        g = torch.Generator().manual_seed(1000 + i)
        q = torch.randn(H, D, generator=g).to(dtype)
        k = torch.randn(L, H, D, generator=g).to(dtype)
        v = torch.randn(L, H, D, generator=g).to(dtype)
        mask = torch.ones(1, L, dtype=torch.bool)
        o = mod.alibi_attention(q.unsqueeze(0), k, v, mask).squeeze(0)   # q's dtype
        out[f"{name}/meta"] = np.array([H, D, L, int(dtype == torch.bfloat16)])
        for key, t in (("q", q), ("k", k), ("v", v), ("o", o)):
            out[f"{name}/{key}"] = t.view(torch.int16).numpy()
        out[f"{name}/slopes"] = mod.get_slopes(H).float().numpy()
    np.savez_compressed(OUT, **out)
    print(f"wrote {OUT} ({OUT.stat().st_size} bytes)")


if __name__ == "__main__":
    main()
