"""Per-CTA phase timestamps of one decode launch (needs the trace build:
`python flashinfer_b200/build.py --trace`, then FI_B200_LIB=flashinfer_b200/libfi_b200_trace.so).
python scripts/trace_probe.py [bf16|fp8] [kv_len] [hkv] [bs]

Slots (ns on %globaltimer, relative to the earliest CTA entry):
 0 entry  1 work list cached  2 PDL wait passed  3 producer: first page index known
 4 producer: first tile issued  5 consumer: first tile landed  6 consumer: last tile consumed
 7 o / partial written  8 segment loop done  9 fused merge done (last arriver only)"""
import ctypes as C
import os
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
os.environ.setdefault("FI_B200_LIB", str(ROOT / "flashinfer_b200" / "libfi_b200_trace.so"))
sys.path.insert(0, str(ROOT))
import flashinfer_b200 as fi  # noqa: E402
from flashinfer_b200 import _lib  # noqa: E402

DEV = "cuda:0"
NAMES = ["entry", "seg cached", "pdl passed", "page idx", "tile issued", "tile landed",
         "last tile done", "o written", "loop done", "merge done", "bar1 passed", "bar2 passed"]


def main():
    kind = sys.argv[1] if len(sys.argv) > 1 else "bf16"
    kv_len = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
    hkv = int(sys.argv[3]) if len(sys.argv) > 3 else 1
    bs = int(sys.argv[4]) if len(sys.argv) > 4 else 64
    page, d = 16, 128
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
    cache = torch.randn(total, 2, page, hkv, d, device=DEV, dtype=torch.bfloat16)
    cache = cache.to(kv_dtype) if kind != "bf16" else cache
    order = os.environ.get("PAGE_ORDER", "random")
    indices = (torch.randperm(total, device=DEV) if order == "random" else torch.arange(total, device=DEV)).to(torch.int32)
    indptr = torch.tensor([0] + list(np.cumsum(pages)), dtype=torch.int32)
    last = torch.tensor([(kv_len - 1) % page + 1] * bs, dtype=torch.int32)
    q = torch.randn(bs, hq, d, device=DEV, dtype=torch.bfloat16)
    w.plan(indptr, indices, last, hq, hkv, d, page, q_data_type=torch.bfloat16, kv_data_type=kv_dtype)
    ncta = int(w._plan_info[1])
    buf = np.zeros((ncta, 16), dtype=np.uint64)
    for it in range(4):
        flush.zero_()
        torch.cuda.synchronize()
        lib.fi_debug_trace(buf.ctypes.data, ncta)  # clears the device buffer
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        flush.zero_()
        e0.record()
        w.run(q, cache)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3
        lib.fi_debug_trace(buf.ctypes.data, ncta)
    t = buf.astype(np.float64)
    t[t == 0] = np.nan
    t0 = np.nanmin(t[:, 0])
    rel = (t - t0) / 1e3
    print(f"{kind} kv_len={kv_len} hkv={hkv}: {ncta} CTAs, event time {us:.1f} us, "
          f"kernel span {np.nanmax(rel):.1f} us")
    print(f"{'slot':16s} {'n':>4s} {'min':>7s} {'median':>7s} {'p90':>7s} {'max':>7s}   (us since first CTA entry)")
    for i, n in enumerate(NAMES):
        c = rel[:, i]
        c = c[~np.isnan(c)]
        if len(c) == 0:
            continue
        print(f"{i} {n:14s} {len(c):4d} {c.min():7.2f} {np.median(c):7.2f} {np.percentile(c, 90):7.2f} {c.max():7.2f}")
    # where does the spread of the streaming phase come from?  (Hkv = 1, static equal split:
    # CTA i owns pages [i n, (i + 1) n) of the flattened page list)
    dur = rel[:, 6] - rel[:, 5]
    ok = ~np.isnan(dur)
    if hkv == 1 and ok.sum() > 8:
        idx = indices.cpu().numpy().astype(np.float64)
        n = total // ncta
        per = idx[: n * ncta].reshape(ncta, n)
        upper = (per >= total / 2).mean(axis=1)
        smid = buf[:, 15].astype(np.float64) - 1
        print(f"page order {order}: streaming phase {np.nanmin(dur):.1f} .. {np.nanmax(dur):.1f} us, std {np.nanstd(dur):.2f}")
        print(f"  corr(duration, share of pages in the upper half of the cache) = {np.corrcoef(dur[ok], upper[ok])[0, 1]:+.2f}")
        print(f"  corr(duration, mean page index)                              = {np.corrcoef(dur[ok], per.mean(axis=1)[ok])[0, 1]:+.2f}")
        print(f"  corr(duration, sm id)                                        = {np.corrcoef(dur[ok], smid[ok])[0, 1]:+.2f}")
        print(f"  corr(duration, sm id parity / cta slot)                      = {np.corrcoef(dur[ok], (np.arange(ncta) >= ncta // 2)[ok])[0, 1]:+.2f}")
        for lo in range(0, 160, 20):
            m = ok & (smid >= lo) & (smid < lo + 20)
            if m.sum():
                print(f"    SMs {lo:3d}-{lo + 19:3d}: mean duration {dur[m].mean():6.2f} us  (n={m.sum()})")
    # does a CTA that has an SM to itself finish earlier than one that shares its SM?
    smid_all = buf[:, 15].astype(np.int64) - 1
    act = ~np.isnan(rel[:, 6])
    if act.sum() > 8:
        per_sm = np.bincount(smid_all[act & (smid_all >= 0)], minlength=160)
        share = per_sm[np.clip(smid_all, 0, 159)]
        for n in (1, 2):
            m = act & (share == n)
            if m.sum():
                print(f"CTAs on an SM with {n} active CTA(s): n={m.sum():3d}  last tile done median {np.median(rel[m, 6]):7.2f} "
                      f"min {rel[m, 6].min():7.2f} max {rel[m, 6].max():7.2f} us")
    print("per-CTA deltas (median / max, us):")
    for a, b in [(0, 1), (1, 2), (2, 3), (3, 4), (4, 5), (5, 6), (6, 10), (10, 11), (11, 7), (6, 7), (7, 8), (7, 9)]:
        dlt = rel[:, b] - rel[:, a]
        dlt = dlt[~np.isnan(dlt)]
        if len(dlt):
            print(f"  {NAMES[a]:>14s} -> {NAMES[b]:14s}: {np.median(dlt):6.2f} / {dlt.max():6.2f}")


if __name__ == "__main__":
    main()
