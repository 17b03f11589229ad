"""Plain decode at GQA groups 8 / 12 / 16 (Llama-3-405B: 128 / 8 heads) and their 8-way per-rank shapes, bf16, C2-like
batch: groups of 9 .. 16 run as one 16-column unit (FI_PLAN_WIDE_UNITS).  python scripts/gqa_group_probe.py [hq hkv]"""
import sys, numpy as np, torch
sys.path.insert(0, str(__import__("pathlib").Path(__file__).resolve().parent.parent))
import flashinfer_b200 as fi
DEV="cuda:0"
def one(hq, hkv, bs, kv_len, kv_dtype=torch.bfloat16):
    d, page = 128, 16
    pages=[kv_len//page]*bs; total=sum(pages)
    cache=torch.randn(total,2,page,hkv,d,device=DEV,dtype=torch.bfloat16).to(kv_dtype)
    indices=torch.randperm(total).to(torch.int32)
    indptr=torch.tensor([0]+list(np.cumsum(pages)),dtype=torch.int32)
    last=torch.full((bs,),page,dtype=torch.int32)
    w=fi.BatchDecodeWithPagedKVCacheWrapper(torch.zeros(128<<20,dtype=torch.uint8,device=DEV),"NHD")
    w.plan(indptr,indices,last,hq,hkv,d,page,q_data_type=torch.bfloat16,kv_data_type=kv_dtype)
    q=torch.randn(bs,hq,d,device=DEV,dtype=torch.bfloat16); out=torch.empty_like(q)
    for _ in range(5): w.run(q,cache,out=out)
    a,b=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); a.record()
    for _ in range(30): w.run(q,cache,out=out)
    b.record(); torch.cuda.synchronize()
    print(f"{str(kv_dtype)[6:]:14s} {hq}/{hkv} heads bs {bs} kv {kv_len}: {a.elapsed_time(b)/30*1e3:.1f} us  (group_size {hq//int(w._plan_info[13])}, virtual kv heads {int(w._plan_info[13])})")
if len(sys.argv) > 2:
    one(int(sys.argv[1]), int(sys.argv[2]), 64, 4096)
else:
    one(64,8,64,4096); one(128,8,64,4096); one(96,8,64,4096); one(16,1,64,4096); one(8,1,64,4096)
    for hq, hkv in ((64, 8), (128, 8), (16, 1)):
        one(hq, hkv, 64, 8192, torch.float8_e4m3fn)
