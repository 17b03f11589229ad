"""cProfile of the host side of plan() + run() at the C2 shape (where do the ~40 + ~25 us go?)."""
import cProfile
import pstats
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import flashinfer_b200 as fi  # noqa: E402
from bench import HEAD_DIM, PAGE, page_table, workload  # noqa: E402

DEV = torch.device("cuda", 0)
wl = workload("c2")
kv_lens, hq, hkv = wl["kv_lens"], wl["hq"], wl["hkv"]
indptr, indices, last, total = page_table(kv_lens)
indptr, indices, last = indptr.pin_memory(), indices.pin_memory(), last.pin_memory()
cache = torch.randn(total, 2, PAGE, hkv, HEAD_DIM, device=DEV, dtype=torch.bfloat16)
q = torch.randn(len(kv_lens), hq, HEAD_DIM, device=DEV, dtype=torch.bfloat16)
out = torch.empty_like(q)
w = fi.BatchDecodeWithPagedKVCacheWrapper(torch.zeros(128 << 20, dtype=torch.uint8, device=DEV), "NHD")


def plan():
    w.plan(indptr, indices, last, hq, hkv, HEAD_DIM, PAGE, q_data_type=torch.bfloat16)


def run():
    w.run(q, cache, out=out)


for fn in (plan, run):
    for _ in range(20):
        fn()
    torch.cuda.synchronize()
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(300):
        fn()
    pr.disable()
    torch.cuda.synchronize()
    print(f"==== {fn.__name__}: per call = cumtime / 300")
    pstats.Stats(pr).sort_stats("tottime").print_stats(14)
