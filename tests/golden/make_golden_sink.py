"""Generate tests/golden/sink_golden.npz by EXECUTING the reference's own attention-sink oracle
(/root/reference/tests/test_helpers/sink_attention_reference.py: `sink_attention_unified` in its
"incremental" = decode mode, built on `sink_softmax` :23-36; the function the reference's
tests/attention/test_attention_sink*.py compare the kernels against).
Build container only (reads /root/reference):  python tests/golden/make_golden_sink.py

The module is pure PyTorch + einops and is loaded by path, unchanged.  Inputs are stored with the outputs.
"""
import importlib.util
from pathlib import Path

import numpy as np
import torch

REF = Path("/root/reference/tests/test_helpers/sink_attention_reference.py")
OUT = Path(__file__).resolve().parent / "sink_golden.npz"

# (name, batch, kv_len, Hq, Hkv, D, window_left, dtype)
CASES = [
    ("b2_len1_h8x8", 2, 1, 8, 8, 128, -1, torch.float16),
    ("b3_len97_h8x2", 3, 97, 8, 2, 128, -1, torch.float16),
    ("b2_len120_h16x1_bf16", 2, 120, 16, 1, 128, -1, torch.bfloat16),
    ("b2_len120_h8x4_win63", 2, 120, 8, 4, 64, 63, torch.float16),
]


def main():
    spec = importlib.util.spec_from_file_location("sink_attention_reference", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    out = {"names": np.array([c[0] for c in CASES])}
    for i, (name, B, L, Hq, Hkv, D, win, dtype) in enumerate(CASES):
        g = torch.Generator().manual_seed(2000 + i)
        q = torch.randn(B, Hq, D, generator=g).to(dtype)
        k = torch.randn(B, L, Hkv, D, generator=g).to(dtype)
        v = torch.randn(B, L, Hkv, D, generator=g).to(dtype)
        sink = (torch.randn(Hq, generator=g) * 2 + 1).float()
        sm_scale = 1.0 / D ** 0.5
        o = mod.sink_attention_unified(q, k, v, sink, win, True, sm_scale, mode="incremental")
        out[f"{name}/meta"] = np.array([B, L, Hq, Hkv, D, win, int(dtype == torch.bfloat16)])
        for key, t in (("q", q), ("k", k), ("v", v), ("o", o.to(dtype))):
            out[f"{name}/{key}"] = t.contiguous().view(torch.int16).numpy()
        out[f"{name}/sink"] = sink.numpy()
    np.savez_compressed(OUT, **out)
    print(f"wrote {OUT} ({OUT.stat().st_size} bytes)")


if __name__ == "__main__":
    main()
