"""Where does the end-to-end step (plan + q H2D + run + o D2H) spend its time?  C2 shape.
python scripts/e2e_probe.py [pinned|pageable]"""
import sys
import time
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import flashinfer_b200 as fi  # noqa: E402
from bench import HEAD_DIM, PAGE, page_table, workload  # noqa: E402

DEV = torch.device("cuda", 0)


def main():
    pinned = (sys.argv[1] if len(sys.argv) > 1 else "pinned") == "pinned"
    wl = workload("c2")
    kv_lens, hq, hkv = wl["kv_lens"], wl["hq"], wl["hkv"]
    bs = len(kv_lens)
    indptr, indices, last, total = page_table(kv_lens)
    if pinned:
        indptr, indices, last = indptr.pin_memory(), indices.pin_memory(), last.pin_memory()
    cache = torch.randn(total, 2, PAGE, hkv, HEAD_DIM, device=DEV, dtype=torch.bfloat16)
    q_host = torch.randn(bs, hq, HEAD_DIM).to(torch.bfloat16).pin_memory()
    o_host = torch.empty_like(q_host).pin_memory()
    q = q_host.to(DEV)
    out = torch.empty_like(q)
    ws = torch.zeros(128 << 20, dtype=torch.uint8, device=DEV)
    w = fi.BatchDecodeWithPagedKVCacheWrapper(ws, "NHD")

    def plan():
        w.plan(indptr, indices, last, hq, hkv, HEAD_DIM, PAGE, q_data_type=torch.bfloat16)

    def t(fn, n=30, sync_each=True):
        ts = []
        for _ in range(n):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            fn()
            if sync_each:
                torch.cuda.synchronize()
            ts.append((time.perf_counter() - t0) * 1e6)
        return float(np.median(ts[3:]))

    plan()
    w.run(q, cache, out=out)
    print(f"page table in {'pinned' if pinned else 'pageable'} host memory")
    print(f"plan() host call only      : {t(plan, sync_each=False):7.1f} us")
    print(f"plan() + sync              : {t(plan):7.1f} us")
    print(f"q H2D + sync               : {t(lambda: q.copy_(q_host, non_blocking=True)):7.1f} us")
    print(f"run() host call only       : {t(lambda: w.run(q, cache, out=out), sync_each=False):7.1f} us")
    print(f"run() + sync               : {t(lambda: w.run(q, cache, out=out)):7.1f} us")
    print(f"o D2H + sync               : {t(lambda: o_host.copy_(out, non_blocking=True)):7.1f} us")

    def e2e():
        plan()
        q.copy_(q_host, non_blocking=True)
        w.run(q, cache, out=out)
        o_host.copy_(out, non_blocking=True)

    print(f"plan + H2D + run + D2H     : {t(e2e):7.1f} us")


if __name__ == "__main__":
    main()
