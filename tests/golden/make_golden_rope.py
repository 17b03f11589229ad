"""Generate tests/golden/rope_golden.npz by EXECUTING the reference's own pure-PyTorch helpers:
RoPE from /root/reference/tests/test_helpers/rope_reference.py (`generate_cos_sin_f32_cache`,
`apply_rotary_pos_emb` / `rotate_half`: LLaMA "rotate half" convention, the one
include/flashinfer/pos_enc.cuh:79-100 implements) followed by the reference test oracle
`attention_ref` (tests/attention/test_blackwell_fmha.py:11-54, lifted as in make_golden.py).
The decode kernel's contract (decode.cuh:445-466,72-76): q rotated at position kv_len - 1, K row j
at position j, V untouched.  Build container only:  python tests/golden/make_golden_rope.py
"""
import importlib.util
import math
from pathlib import Path

import numpy as np
import torch

from make_golden import load_reference_attention_ref

REF = Path("/root/reference/tests/test_helpers/rope_reference.py")
OUT = Path(__file__).resolve().parent / "rope_golden.npz"

# (name, Hq, Hkv, D, kv_len, dtype, rope_theta)
CASES = [
    ("g1_d128_len1", 4, 4, 128, 1, torch.float16, 1e4),
    ("g4_d128_len54", 8, 2, 128, 54, torch.float16, 1e4),
    ("g4_d128_len97_bf16", 8, 2, 128, 97, torch.bfloat16, 1e4),
    ("g8_d128_len300_theta5e5", 8, 1, 128, 300, torch.float16, 5e5),
    ("g2_d64_len130", 4, 2, 64, 130, torch.float16, 1e4),
    ("g2_d256_len65", 4, 2, 256, 65, torch.float16, 1e4),
]


def main():
    spec = importlib.util.spec_from_file_location("rope_reference", REF)
    rr = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(rr)
    attention_ref = load_reference_attention_ref()
    out = {"names": np.array([c[0] for c in CASES])}
    for i, (name, Hq, Hkv, D, L, dtype, theta) in enumerate(CASES):
        g = torch.Generator().manual_seed(2000 + i)
        q = torch.randn(Hq, D, generator=g).to(dtype)
        k = torch.randn(L, Hkv, D, generator=g).to(dtype)
        v = torch.randn(L, Hkv, D, generator=g).to(dtype)
        cos, sin = rr.generate_cos_sin_f32_cache(L, D, theta=theta, device="cpu")   # [L, D] f32
        # K row j at position j; q at position L - 1 (fp32 in, fp32 out: no rounding of the rotated values)
        _, k_rot = rr.apply_rotary_pos_emb(k.float(), k.float(), cos, sin)
        q_rot, _ = rr.apply_rotary_pos_emb(q.float().unsqueeze(0), q.float().unsqueeze(0), cos[L - 1:L], sin[L - 1:L])
        grp = Hq // Hkv
        o, lse2 = attention_ref(1, q_rot, k_rot.repeat_interleave(grp, dim=1),
                                v.float().repeat_interleave(grp, dim=1), False, 1.0 / math.sqrt(D))
        out[f"{name}/meta"] = np.array([Hq, Hkv, D, L, int(dtype == torch.bfloat16)])
        out[f"{name}/theta"] = np.array([theta], dtype=np.float64)
        for key, t in (("q", q), ("k", k), ("v", v)):
            out[f"{name}/{key}"] = t.view(torch.int16).numpy()
        out[f"{name}/o"] = o.float().numpy().reshape(Hq, D)
        out[f"{name}/lse2"] = lse2.float().numpy().reshape(Hq)
    np.savez_compressed(OUT, **out)
    print(f"wrote {OUT} ({OUT.stat().st_size} bytes)")


if __name__ == "__main__":
    main()
