"""Developer probe: time run() for one config under several tuning overrides.
python scripts/perf_probe.py [c2|c3|c3r|c4|c5r] (r = per-rank shape of the 8-GPU split)."""
import sys
import time
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import flashinfer_b200 as fi  # noqa: E402
from flashinfer_b200 import _lib  # noqa: E402

DEV = "cuda:0"


def config(name):
    if name == "c2":
        return dict(kv_lens=[4096] * 64, hq=32, hkv=8, kv_dtype=torch.bfloat16)
    if name == "c2r":
        return dict(kv_lens=[4096] * 64, hq=4, hkv=1, kv_dtype=torch.bfloat16)
    if name == "c3":
        return dict(kv_lens=[8192] * 128, hq=64, hkv=8, kv_dtype=torch.float8_e4m3fn)
    if name == "c3r":
        return dict(kv_lens=[8192] * 128, hq=8, hkv=1, kv_dtype=torch.float8_e4m3fn)
    if name == "c4":
        rng = np.random.default_rng(0)
        lens = np.exp(rng.uniform(np.log(512), np.log(32768), 256)).round().astype(int).tolist()
        return dict(kv_lens=lens, hq=32, hkv=8, kv_dtype=torch.bfloat16)
    if name == "c5r":
        return dict(kv_lens=[131072] * 16, hq=8, hkv=1, kv_dtype=torch.bfloat16)
    if name == "c70":
        return dict(kv_lens=[4096] * 64, hq=64, hkv=8, kv_dtype=torch.bfloat16)
    raise SystemExit(f"unknown config {name}")


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "c2"
    layout = sys.argv[2] if len(sys.argv) > 2 else "NHD"
    c = config(name)
    page, d = 16, 128
    kv_lens, hq, hkv, kv_dtype = c["kv_lens"], c["hq"], c["hkv"], c["kv_dtype"]
    pages = [-(-l // page) for l in kv_lens]
    total = sum(pages)
    shape = (total, 2, page, hkv, d) if layout == "NHD" else (total, 2, hkv, page, d)
    es = 1 if kv_dtype == torch.float8_e4m3fn else 2
    cache = torch.randn(shape, device=DEV, dtype=torch.bfloat16)
    cache = cache.to(kv_dtype) if es == 1 else cache
    indices = torch.randperm(total, device=DEV).to(torch.int32)
    indptr = torch.tensor([0] + list(np.cumsum(pages)), dtype=torch.int32)
    last = torch.tensor([((l - 1) % page + 1) for l in kv_lens], dtype=torch.int32)
    q = torch.randn(len(kv_lens), hq, d, device=DEV, dtype=torch.bfloat16)
    ws = torch.zeros(128 << 20, dtype=torch.uint8, device=DEV)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)
    nbytes = sum(kv_lens) * hkv * d * 2 * es + len(kv_lens) * hq * d * 4
    lib = _lib.load()
    print(f"{name} {layout}: {nbytes/1e6:.1f} MB algorithmic; ideal {nbytes/8e12*1e6:.1f} us @8TB/s")
    w = fi.BatchDecodeWithPagedKVCacheWrapper(ws, layout)
    for stages, cps, loader, mt in [(0, 0, 0, 0), (0, 1, 0, 0), (0, 2, 0, 0), (2, 2, 0, 0), (0, 0, 2, 0),
                                    (0, 1, 2, 0), (4, 1, 0, 0), (3, 1, 0, 0)]:
        lib.fi_set_tuning(stages, cps, loader, mt)
        try:
            t0 = time.perf_counter()
            w.plan(indptr, indices, last, hq, hkv, d, page, q_data_type=torch.bfloat16,
                   kv_data_type=kv_dtype)
            plan_ms = (time.perf_counter() - t0) * 1e3
            for _ in range(3):
                w.run(q, cache)
            torch.cuda.synchronize()
            ts = []
            for _ in range(20):
                flush.zero_()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                w.run(q, cache)
                e1.record()
                torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1) * 1e3)
            ts = np.array(ts)
            info = w._plan_info
            print(f"stages={stages} cps={cps} loader={loader}: used stages={int(info[22])} grid={int(info[1])} "
                  f"merge={int(info[3])} | median {np.median(ts):.1f} us min {ts.min():.1f} us -> "
                  f"{nbytes/np.median(ts)/1e6:.2f} TB/s (loader used {lib.fi_last_loader()}) plan {plan_ms:.2f} ms")
        except Exception as e:  # noqa: BLE001
            print(f"stages={stages} cps={cps} loader={loader}: FAILED {e}")
    lib.fi_set_tuning(0, 0, 0, 0)


if __name__ == "__main__":
    main()
