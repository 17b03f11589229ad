"""Shared test helpers: golden fixtures, synthetic paged caches, plan replay."""
from __future__ import annotations

import ctypes as C
import hashlib
from pathlib import Path
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

GOLDEN = Path(__file__).resolve().parent / "golden" / "decode_golden.npz"


# ---------------------------------------------------------------- golden fixtures
def golden_cases():
    z = np.load(GOLDEN)
    return z, [str(n) for n in z["names"]]


def golden_inputs(z, name):
    Hq, Hkv, D, L, is_bf16, seed, stored = [int(x) for x in z[f"{name}/meta"]]
    dtype = torch.bfloat16 if is_bf16 else torch.float16
    if stored:
        q = torch.from_numpy(z[f"{name}/q"].copy()).view(dtype)
        k = torch.from_numpy(z[f"{name}/k"].copy()).view(dtype)
        v = torch.from_numpy(z[f"{name}/v"].copy()).view(dtype)
    else:
        g = torch.Generator().manual_seed(seed)
        q = torch.randn(Hq, D, generator=g).to(dtype)
        k = torch.randn(L, Hkv, D, generator=g).to(dtype)
        v = torch.randn(L, Hkv, D, generator=g).to(dtype)
    h = hashlib.sha256()
    for t in (q, k, v):
        h.update(t.view(torch.int16).numpy().tobytes())
    ok = h.digest() == z[f"{name}/hash"].tobytes()
    return q, k, v, ok


# ---------------------------------------------------------------- synthetic paged caches
def make_paged_case(
    kv_lens: Sequence[int],
    num_kv_heads: int,
    head_dim: int,
    page_size: int,
    kv_layout: str = "NHD",
    dtype: torch.dtype = torch.float16,
    seed: int = 0,
    extra_pages: int = 3,
    shuffle: bool = True,
    device: str = "cpu",
):
    """Random 5-D paged cache with a shuffled page table (values ~ N(0,1)
    generated in fp32 then cast, test_batch_decode_kernels.py:109-110)."""
    g = torch.Generator().manual_seed(seed)
    pages_per = [(-(-l // page_size)) for l in kv_lens]
    total = sum(pages_per)
    num_pages = total + extra_pages
    if kv_layout == "NHD":
        shape = (num_pages, 2, page_size, num_kv_heads, head_dim)
    else:
        shape = (num_pages, 2, num_kv_heads, page_size, head_dim)
    cache32 = torch.randn(*shape, generator=g)
    perm = torch.randperm(num_pages, generator=g) if shuffle else torch.arange(num_pages)
    indices = perm[:total].to(torch.int32)
    indptr = torch.zeros(len(kv_lens) + 1, dtype=torch.int32)
    indptr[1:] = torch.cumsum(torch.tensor(pages_per, dtype=torch.int32), 0)
    last = torch.tensor([((l - 1) % page_size + 1) if l > 0 else 0 for l in kv_lens], dtype=torch.int32)
    if dtype in (torch.float8_e4m3fn, torch.float8_e5m2):
        cache = cache32.to(dtype)
    else:
        cache = cache32.to(dtype)
    return cache.to(device), indptr, indices, last


def make_q(bs, num_qo_heads, head_dim, dtype, seed=1, device="cpu"):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(bs, num_qo_heads, head_dim, generator=g).to(dtype).to(device)


# ---------------------------------------------------------------- plan through the C-ABI (host only)
class Segment(C.Structure):
    _fields_ = [("request", C.c_int32), ("kv_head", C.c_int32), ("tile_begin", C.c_int32),
                ("tile_end", C.c_int32), ("slot", C.c_int32), ("tok_lo", C.c_int32),
                ("tok_hi", C.c_int32), ("pad", C.c_int32)]


class MergeUnit(C.Structure):
    _fields_ = [("request", C.c_int32), ("kv_head", C.c_int32), ("slot_begin", C.c_int32),
                ("count", C.c_int32)]


# plan-info slots (flashinfer_b200/csrc/decode_params.h PlanInfoSlot)
PI = dict(magic=0, num_ctas=1, num_segments=2, num_merge=3, num_slots=4, off_header=5,
          num_items=6, off_segments=7, off_merge=8, off_tmp_o=9, off_tmp_lse=10,
          batch_size=11, hq=12, hkv=13, num_kv_heads=13, head_dim=14, page_size=15, q_dtype=16, kv_dtype=17,
          cuda_graph=18, merge_grid=19, total_tiles=20, window_left=21, stages=22, int_bytes=23,
          off_counters=24, off_item_heads=25, off_work_counter=26, kv_head_div=27, off_indices=28,
          num_indices=29, indices_stable=30, head_pairs=31)
SEG_BYTES = 48    # sizeof(Segment)
HEAD_BYTES = 224  # sizeof(ItemHead)
INLINE_PAGES = 16  # kItemInlinePages
INLINE_SEGS = 3   # kItemInlineSegs
TILE = 64        # tokens per tile for 16-bit KV (csrc/decode_params.h tile_tokens)
TILE_FP8 = 128   # ... and for fp8 KV


def host_plan(indptr, last_page_len, num_qo_heads, num_kv_heads, head_dim, page_size,
              num_ctas, q_dtype=1, kv_dtype=1, window_left=-1, cuda_graph=0,
              kv_len=None, int_bytes=8 << 20, float_bytes=128 << 20, fixed_split_pages=0,
              disable_split_kv=0, indices=None, flags=0):
    """Run fi_batch_decode_plan in host-only mode (device = -1) and decode the work list."""
    from flashinfer_b200 import _lib

    lib = _lib.load()
    ws = np.zeros(int_bytes, dtype=np.uint8)
    ip = np.ascontiguousarray(np.asarray(indptr, dtype=np.int32))
    lp = np.ascontiguousarray(np.asarray(last_page_len, dtype=np.int32))
    kl = None if kv_len is None else np.ascontiguousarray(np.asarray(kv_len, dtype=np.int32))
    info = np.zeros(lib.fi_plan_info_len(), dtype=np.int64)
    ix = None if indices is None else np.ascontiguousarray(np.asarray(indices, dtype=np.int32))
    if flags:  # FI_PLAN_* bits exist in the argument-block form only
        a = _lib.DecodePlanArgs()
        a.float_ws, a.float_ws_bytes = None, float_bytes
        a.int_ws = a.pinned_int_ws = ws.ctypes.data
        a.int_ws_bytes = int_bytes
        a.indptr_host, a.last_page_len_host = ip.ctypes.data, lp.ctypes.data
        a.kv_len_host = None if kl is None else kl.ctypes.data
        a.batch_size, a.num_qo_heads, a.num_kv_heads = len(lp), num_qo_heads, num_kv_heads
        a.head_dim, a.page_size, a.q_dtype, a.kv_dtype = head_dim, page_size, q_dtype, kv_dtype
        a.window_left, a.cuda_graph = window_left, cuda_graph
        a.fixed_split_pages, a.disable_split_kv = fixed_split_pages, disable_split_kv
        a.device, a.num_ctas_override, a.flags, a.stream = -1, num_ctas, flags, None
        a.kv_indices_host = None if indices is None else ix.ctypes.data
        a.num_indices = 0 if indices is None else len(ix)
        a.plan_info_out = info.ctypes.data
        rc = lib.fi_batch_decode_plan_args(C.byref(a))
    else:
        rc = lib.fi_batch_decode_plan(
            None, float_bytes, ws.ctypes.data, ws.ctypes.data, int_bytes, ip.ctypes.data,
            lp.ctypes.data, None if kl is None else kl.ctypes.data, len(lp), num_qo_heads,
            num_kv_heads, head_dim, page_size, q_dtype, kv_dtype, window_left, cuda_graph,
            fixed_split_pages, disable_split_kv, -1, num_ctas, None,
            None if indices is None else ix.ctypes.data, 0 if indices is None else len(ix), info.ctypes.data)
    if rc != 0:
        return rc, _lib.last_error(), None
    hdr = ws[info[PI["off_header"]]:info[PI["off_header"]] + 32].view(np.int32)
    ncta = int(info[PI["num_ctas"]])
    nitems = int(info[PI["num_items"]])
    nseg = int(hdr[1])
    assert hdr[0] == nitems and nseg == info[PI["num_segments"]]
    # Segment records are 48 bytes (12 int32): 8 work fields, page_base, 3 pad
    seg_full = ws[info[PI["off_segments"]]:info[PI["off_segments"]] + SEG_BYTES * nseg].view(np.int32)
    seg_full = seg_full.reshape(nseg, SEG_BYTES // 4).copy()
    segs = seg_full[:, :8].copy()
    nm = int(hdr[2])
    merges = ws[info[PI["off_merge"]]:info[PI["off_merge"]] + 16 * nm].view(np.int32).reshape(nm, 4).copy()
    # item heads (224 bytes): first, count, num_items, num_pages, kItemInlineSegs inline Segment
    # records, then kItemInlinePages page ids; one head per item, padded with empty heads up to the grid
    nheads = max(nitems, ncta)
    heads = ws[info[PI["off_item_heads"]]:info[PI["off_item_heads"]] + HEAD_BYTES * nheads].view(np.int32)
    heads = heads.reshape(nheads, HEAD_BYTES // 4).copy()
    # the segment list starts right after the last head (the kernel relies on it)
    assert info[PI["off_segments"]] == info[PI["off_item_heads"]] + HEAD_BYTES * nheads
    item_indptr = np.concatenate([heads[:nitems, 0], [nseg]]).astype(np.int32)
    # structural checks that need the raw records
    assert np.array_equal(seg_full[:, 8], ip[segs[:, 0]]), "page_base != kv_indptr[request]"
    for r in seg_full:  # partial segments carry their unit's (count, slot_begin)
        if r[7] >= 0:
            assert (r[9], r[10]) == (merges[r[7], 3], merges[r[7], 2]), "merge_count / merge_slot"
            assert merges[r[7], 2] <= r[4] < merges[r[7], 2] + merges[r[7], 3]
    assert np.array_equal(heads[:nitems, 1], np.diff(item_indptr))
    assert np.all(heads[:, 2] == nitems)
    assert np.all(heads[nitems:, 1] == 0), "heads past the last item must be empty"
    for c in range(nitems):
        for k in range(min(int(heads[c, 1]), INLINE_SEGS)):
            inl = heads[c, 4 + 12 * k:4 + 12 * (k + 1)]
            assert np.array_equal(inl, seg_full[heads[c, 0] + k]), "inline segment != list segment"
    # inline page ids: exactly the table entries the item's first tiles gather (or none at all)
    tile = TILE if kv_dtype in (0, 1) else TILE_FP8
    if info[PI["head_pairs"]]:
        tile //= 2
    for c in range(nheads):
        npg = int(heads[c, 3])
        if ix is None or c >= nitems or heads[c, 1] == 0:
            assert npg == 0
            continue
        s0 = seg_full[heads[c, 0]]
        if s0[3] <= s0[2]:
            assert npg == 0
            continue
        first_page = int(s0[2]) * tile // page_size
        req_pages = int(ip[s0[0] + 1] - ip[s0[0]])
        assert npg == min(INLINE_PAGES, req_pages - first_page) and npg > 0
        exp = ix[s0[8] + first_page: s0[8] + first_page + npg]
        assert np.array_equal(heads[c, 4 + 12 * INLINE_SEGS: 4 + 12 * INLINE_SEGS + npg], exp), "inline pages"
    if ix is not None and not cuda_graph:
        oi = int(info[PI["off_indices"]])
        assert oi >= info[PI["off_segments"]] + SEG_BYTES * nseg and info[PI["num_indices"]] == len(ix)
        assert np.array_equal(ws[oi:oi + 4 * len(ix)].view(np.int32), ix), "workspace copy of kv_indices"
        assert info[PI["indices_stable"]] == 1
    else:
        assert info[PI["off_indices"]] == -1
    wc = ws[info[PI["off_work_counter"]]:info[PI["off_work_counter"]] + 4].view(np.int32)
    cnt = ws[info[PI["off_counters"]]:info[PI["off_counters"]] + 4 * max(nm, 1)].view(np.int32)
    assert wc[0] == 0 and np.all(cnt == 0), "scheduler / merge counters must start at zero"
    return 0, "", dict(info=info, header=hdr.copy(), item_indptr=item_indptr, segments=segs,
                       merges=merges, item_heads=heads)


def replay_plan(plan, kv_lens, num_kv_heads, window_left=-1, tile=TILE, chunked=False):
    """Pure-Python replay of the work list: returns per-(request, head) token visit counts
    and checks the structural invariants.  Raises AssertionError on any inconsistency."""
    segs, merges, cta_indptr, hdr = plan["segments"], plan["merges"], plan["item_indptr"], plan["header"]
    bs = len(kv_lens)
    visits = [np.zeros(max(l, 0), dtype=np.int32) for l in kv_lens for _ in range(num_kv_heads)]
    direct = np.zeros((bs, num_kv_heads), dtype=np.int32)
    slot_owner = {}
    assert cta_indptr[0] == 0 and np.all(np.diff(cta_indptr) >= 0)
    assert cta_indptr[-1] == len(segs) == hdr[1]
    for r, h, tb, te, slot, lo, hi, _ in segs:
        assert 0 <= r < bs and 0 <= h < num_kv_heads
        L = kv_lens[r]
        exp_lo = max(0, L - 1 - window_left) if window_left >= 0 else 0
        if chunked:  # fixed_split_size: a segment attends its own chunk of the request
            assert (exp_lo <= lo < hi <= L) or (lo == exp_lo and hi == L == 0) or L == 0, (lo, hi, exp_lo, L)
        else:
            assert lo == exp_lo and hi == L, (lo, hi, exp_lo, L)
        assert tb <= te
        a, b = max(tb * tile, lo), min(te * tile, hi)
        if b > a:
            visits[r * num_kv_heads + h][a:b] += 1
        if slot < 0:
            direct[r, h] += 1
        else:
            assert slot not in slot_owner, "partial slot written twice"
            slot_owner[slot] = (r, h)
    merged = np.zeros((bs, num_kv_heads), dtype=np.int32)
    for r, h, sb, cnt in merges:
        assert cnt >= 2, "a merge unit must have at least two partials"
        merged[r, h] += 1
        for s in range(sb, sb + cnt):
            assert slot_owner.get(s) == (r, h), "merge unit reads a slot it does not own"
            del slot_owner[s]
    assert not slot_owner, "partial slots never merged"
    # every unit is written exactly once: directly or through exactly one merge
    assert np.all(direct + merged == 1), "unit written != once"
    for r in range(bs):
        L = kv_lens[r]
        lo = max(0, L - 1 - window_left) if window_left >= 0 else 0
        for h in range(num_kv_heads):
            v = visits[r * num_kv_heads + h]
            assert np.all(v[lo:] == 1), f"tokens of request {r} head {h} not visited exactly once"
            assert np.all(v[:lo] == 0)
    return visits
