"""Host time of fi_batch_decode_plan_args in host-only mode (device = -1, no GPU needed): C2, C3 (head pairs),
the 8-way per-rank C2 shape and C4, then its scaling with the grid and the batch size.
WITH_IDX=0 leaves the page table out; FI_B200_LIB=<other build> times another build of the library."""
import sys, time, ctypes as C, numpy as np
from pathlib import Path
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
from flashinfer_b200 import _lib
from test_abi_and_plan import lens_to_table
lib = _lib.load()
import os
WITH_IDX = os.environ.get('WITH_IDX','1')=='1'
def bench(kv_lens, hq, hkv, d, page, kvd, flags, n=3000, ctas=296):
    ip, lp = lens_to_table(kv_lens, page)
    ix = np.random.default_rng(0).permutation(int(ip[-1])).astype(np.int32)
    ws = np.zeros(8 << 20, dtype=np.uint8); info = np.zeros(lib.fi_plan_info_len(), dtype=np.int64)
    a = _lib.DecodePlanArgs()
    a.float_ws, a.float_ws_bytes = None, 128 << 20
    a.int_ws = a.pinned_int_ws = ws.ctypes.data; a.int_ws_bytes = 8 << 20
    a.indptr_host, a.last_page_len_host, a.kv_len_host = ip.ctypes.data, lp.ctypes.data, None
    a.batch_size, a.num_qo_heads, a.num_kv_heads = len(lp), hq, hkv
    a.head_dim, a.page_size, a.q_dtype, a.kv_dtype = d, page, 1, kvd
    a.window_left, a.cuda_graph, a.fixed_split_pages, a.disable_split_kv = -1, 0, 0, 0
    a.device, a.num_ctas_override, a.flags, a.stream = -1, ctas, flags, None
    a.kv_indices_host, a.num_indices = (ix.ctypes.data, len(ix)) if WITH_IDX else (None, 0)
    a.plan_info_out = info.ctypes.data
    ref = C.byref(a)
    for _ in range(200): assert lib.fi_batch_decode_plan_args(ref) == 0
    best = 1e9
    for rep in range(5):
        t = time.perf_counter()
        for _ in range(n): lib.fi_batch_decode_plan_args(ref)
        best = min(best, (time.perf_counter() - t) / n * 1e6)
    return best
print("C2  ", round(bench([4096]*64, 32, 8, 128, 16, 1, 0), 2), "us")
print("C3  ", round(bench([8192]*128, 64, 8, 128, 16, _lib.DTYPE_E4M3, _lib.PLAN_HEAD_PAIRS, n=1000), 2), "us")
print("C2/8", round(bench([4096]*64, 4, 1, 128, 16, 1, 0), 2), "us")
rng = np.random.default_rng(1)
print("C4  ", round(bench(list(np.exp(rng.uniform(np.log(512), np.log(32768), 256)).astype(int)), 32, 8, 128, 16, 1, 0, n=300), 2), "us")
print("--- scaling")
for ctas in (8, 64, 296, 1184):
    print("C2 ctas", ctas, round(bench([4096]*64, 32, 8, 128, 16, 1, 0, ctas=ctas), 2))
for bs in (1, 8, 64, 512):
    print("bs", bs, round(bench([4096]*bs, 32, 8, 128, 16, 1, 0), 2))
