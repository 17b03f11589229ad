"""Development aid: fp8 KV decode with a head-pair plan against the same launch without
(FI_B200_HEAD_PAIRS=0 is read at every plan) — per-request, per-head-parity error of o and lse."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import flashinfer_b200 as fi
from helpers import make_paged_case, make_q

DEV = "cuda:0"
hq, hkv, d, page = 64, 8, 128, 16
lens = [int(x) for x in sys.argv[1:]] or [1, 54, 64, 65, 128, 129, 200, 512, 777, 4099, 9000, 20000]
cache, indptr, indices, last = make_paged_case(lens, hkv, d, page, "NHD", torch.float32, seed=95)
cache = (cache * 0.5).to(torch.float8_e4m3fn).to(DEV)
q = make_q(len(lens), hq, d, torch.bfloat16, seed=96).to(DEV)
ws = torch.empty(128 << 20, dtype=torch.uint8, device=DEV)
res = {}
for pairs in ("0", "1"):
    os.environ["FI_B200_HEAD_PAIRS"] = pairs
    w = fi.BatchDecodeWithPagedKVCacheWrapper(ws, "NHD")
    w.plan(indptr, indices, last, hq, hkv, d, page, q_data_type=torch.bfloat16, kv_data_type=cache.dtype)
    print("pairs env", pairs, "plan slot31", int(w._plan_info[31]), "merge units", int(w._plan_info[3]))
    o, lse = w.run(q, cache, return_lse=True)
    torch.cuda.synchronize()
    res[pairs] = (o.float().cpu(), lse.cpu())
do = (res["0"][0] - res["1"][0]).abs().amax(-1)  # [bs, hq]
dl = (res["0"][1] - res["1"][1]).abs()
g = hq // hkv
for i, L in enumerate(lens):
    par = [[], []]
    for h in range(hq):
        par[(h // g) & 1].append((float(do[i, h]), float(dl[i, h])))
    s = "  ".join(f"head%2={k}: o {max(x[0] for x in par[k]):.4f} lse {max(x[1] for x in par[k]):.4f}" for k in (0, 1))
    print(f"len {L:6d}  {s}")
