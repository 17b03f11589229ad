"""Compare the work lists of two builds of the planner byte for byte over 3000 fuzzed argument sets (host-only
mode, no GPU needed): python scripts/plan_ab.py <path to the other libfi_b200.so>.  The whole int workspace
(pre-filled with a pattern) and the plan-info vector must be identical; used to check planner refactorings."""
import sys, os, ctypes as C, numpy as np, random
from pathlib import Path
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
from flashinfer_b200 import _lib
from test_abi_and_plan import lens_to_table
new = _lib.load()
old = C.CDLL(sys.argv[1])
old.fi_batch_decode_plan_args.restype = C.c_int
old.fi_batch_decode_plan_args.argtypes = [C.c_void_p]
def run(lib, c, ws, info):
    ip, lp = lens_to_table(c["kv_lens"], c["page"])
    ix = np.random.default_rng(7).permutation(int(ip[-1]) + 3)[: int(ip[-1])].astype(np.int32) if c["idx"] else None
    a = _lib.DecodePlanArgs()
    a.float_ws, a.float_ws_bytes = None, 1 << 36
    a.int_ws = a.pinned_int_ws = ws.ctypes.data; a.int_ws_bytes = ws.size
    a.indptr_host, a.last_page_len_host, a.kv_len_host = ip.ctypes.data, lp.ctypes.data, None
    a.batch_size, a.num_qo_heads, a.num_kv_heads = len(lp), c["hq"], c["hkv"]
    a.head_dim, a.page_size, a.q_dtype, a.kv_dtype = c["d"], c["page"], 1, c["kvd"]
    a.window_left, a.cuda_graph = c["win"], c["graph"]
    a.fixed_split_pages, a.disable_split_kv = c["fixed"], c["nosplit"]
    a.device, a.num_ctas_override, a.flags, a.stream = -1, c["ctas"], c["flags"], None
    a.kv_indices_host, a.num_indices = (ix.ctypes.data, len(ix)) if ix is not None else (None, 0)
    a.plan_info_out = info.ctypes.data
    return lib.fi_batch_decode_plan_args(C.byref(a) if lib is new else C.addressof(a))
rnd = random.Random(11)
lens_pool = [0, 1, 63, 64, 65, 127, 128, 129, 4096, 8192]
n_same = 0
cases = [dict(kv_lens=[4096]*64, page=16, hkv=8, hq=32, d=128, kvd=1, ctas=296, win=-1, graph=0, fixed=0, nosplit=0, flags=0, idx=True),
         dict(kv_lens=[8192]*128, page=16, hkv=8, hq=64, d=128, kvd=_lib.DTYPE_E4M3, ctas=296, win=-1, graph=0, fixed=0, nosplit=0, flags=3, idx=True),
         dict(kv_lens=[131072]*16, page=16, hkv=8, hq=64, d=128, kvd=1, ctas=296, win=-1, graph=0, fixed=0, nosplit=0, flags=2, idx=True)]
for _ in range(3000):
    bs = rnd.randint(1, 40)
    uni = rnd.random() < 0.4
    L = rnd.choice(lens_pool + [rnd.randint(0, 9000)])
    kv = [L]*bs if uni else [rnd.choice(lens_pool + [rnd.randint(0, 9000)]) for _ in range(bs)]
    hkv = rnd.choice([1, 2, 3, 4, 8]); g = rnd.choice([1, 2, 4, 5, 8, 12, 16, 32])
    mode = rnd.choice(["d", "d", "f", "n"])
    cases.append(dict(kv_lens=kv, page=rnd.choice([1, 2, 4, 5, 8, 16, 32, 64]), hkv=hkv, hq=hkv*g, d=rnd.choice([64, 128, 256, 512]),
                      kvd=rnd.choice([1, _lib.DTYPE_E4M3]), ctas=rnd.choice([1, 7, 31, 148, 296, 600]), win=rnd.choice([-1, -1, 0, 17, 1000]),
                      graph=rnd.randint(0, 1), fixed=rnd.choice([1, 3, 16]) if mode == "f" else 0, nosplit=int(mode == "n"),
                      flags=rnd.randint(0, 3), idx=rnd.random() < 0.6))
for c in cases:
    wa = np.full(32 << 20, 0xAB, dtype=np.uint8); wb = np.full(32 << 20, 0xAB, dtype=np.uint8)
    ia = np.zeros(32, dtype=np.int64); ib = np.zeros(32, dtype=np.int64)
    ra, rb = run(new, c, wa, ia), run(old, c, wb, ib)
    assert ra == rb, (ra, rb, c)
    if ra != 0: continue
    assert np.array_equal(ia, ib), c
    assert np.array_equal(wa, wb), ("work list differs", {k: v for k, v in c.items() if k != "kv_lens"}, c["kv_lens"][:5])
    n_same += 1
print("identical work lists:", n_same, "of", len(cases))
