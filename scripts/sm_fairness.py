"""Is the per-CTA streaming rate a property of the SM?  Static split (FI_B200_SCHED_F0=100), trace
build: per-CTA time from first tile landed to last tile consumed over several launches, correlated
between launches and grouped by %smid.  python scripts/sm_fairness.py [bf16|fp8] [kv_len] [hkv]"""
import ctypes as C
import os
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
os.environ.setdefault("FI_B200_LIB", str(ROOT / "flashinfer_b200" / "libfi_b200_trace.so"))
os.environ.setdefault("FI_B200_SCHED_F0", "100")
sys.path.insert(0, str(ROOT))
import flashinfer_b200 as fi  # noqa: E402
from flashinfer_b200 import _lib  # noqa: E402

DEV = "cuda:0"


def main():
    kind = sys.argv[1] if len(sys.argv) > 1 else "bf16"
    kv_len = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
    hkv = int(sys.argv[3]) if len(sys.argv) > 3 else 8
    bs, page, d = 64, 16, 128
    hq = hkv * (4 if kind == "bf16" else 8)
    kv_dtype = torch.bfloat16 if kind == "bf16" else torch.float8_e4m3fn
    lib = _lib.load()
    lib.fi_debug_trace.restype = C.c_int
    lib.fi_debug_trace.argtypes = [C.c_void_p, C.c_int]
    ws = torch.zeros(128 << 20, dtype=torch.uint8, device=DEV)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)
    w = fi.BatchDecodeWithPagedKVCacheWrapper(ws, "NHD")
    pages = [-(-kv_len // page)] * bs
    total = sum(pages)
    q = torch.randn(bs, hq, d, device=DEV, dtype=torch.bfloat16)
    indptr = torch.tensor([0] + list(np.cumsum(pages)), dtype=torch.int32)
    last = torch.tensor([(kv_len - 1) % page + 1] * bs, dtype=torch.int32)
    runs = []
    for it in range(6):
        # a fresh cache and page table every second launch: is the pattern tied to addresses?
        if it % 2 == 0:
            cache = torch.randn(total, 2, page, hkv, d, device=DEV, dtype=torch.bfloat16)
            cache = cache.to(kv_dtype) if kind != "bf16" else cache
            indices = torch.randperm(total, device=DEV).to(torch.int32)
            w.plan(indptr, indices, last, hq, hkv, d, page, q_data_type=torch.bfloat16,
                   kv_data_type=kv_dtype)
        ncta = int(w._plan_info[1])
        buf = np.zeros((ncta, 16), dtype=np.uint64)
        flush.zero_()
        torch.cuda.synchronize()
        lib.fi_debug_trace(buf.ctypes.data, ncta)
        w.run(q, cache)
        torch.cuda.synchronize()
        lib.fi_debug_trace(buf.ctypes.data, ncta)
        t = buf.astype(np.float64)
        runs.append(((t[:, 6] - t[:, 5]) / 1e3, t[:, 15].astype(int) - 1, (t[:, 6] - t[:, 0].min()) / 1e3))
    print(f"{kind} kv_len={kv_len} hkv={hkv}: stream time per CTA (us), static split")
    for i, (dur, sm, fin) in enumerate(runs):
        print(f" launch {i}: min {dur.min():.1f} median {np.median(dur):.1f} max {dur.max():.1f}; "
              f"finish median {np.median(fin):.1f} max {fin.max():.1f}")
    for i in range(1, len(runs)):
        same_map = np.array_equal(runs[0][1], runs[i][1])
        print(f" corr(launch 0, launch {i}) = {np.corrcoef(runs[0][0], runs[i][0])[0, 1]:.3f}  same blockIdx->SM map: {same_map}")
    dur, sm, _ = runs[-1]
    per_sm = np.zeros(sm.max() + 1)
    for s in range(sm.max() + 1):
        per_sm[s] = dur[sm == s].mean() if np.any(sm == s) else np.nan
    mean_all = np.nanmean(np.stack([np.array([r[0][r[1] == s].mean() if np.any(r[1] == s) else np.nan
                                              for s in range(sm.max() + 1)]) for r in runs]), axis=0)
    print(" mean stream time by SM id (all launches), 16 per row:")
    for r0 in range(0, len(mean_all), 16):
        print("  " + " ".join(f"{x:6.1f}" for x in mean_all[r0:r0 + 16]))
    np.save(ROOT / "gpurun_out" / "sm_fairness.npy", np.stack([np.stack([r[0], r[1], r[2]]) for r in runs]))


if __name__ == "__main__":
    main()
