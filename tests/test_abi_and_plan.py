"""CPU: the C-ABI library loads and exports every symbol include/fi_b200.h declares, and the
host planner (csrc/plan.cu) is self-consistent: every KV token of every (request, kv head)
is visited exactly once, every unit is written exactly once, work is balanced over CTAs
(north_star: "plan metadata must be bit-exact self-consistent")."""
import re
from pathlib import Path

import numpy as np
import pytest

from helpers import PI, TILE, TILE_FP8, host_plan, replay_plan

ROOT = Path(__file__).resolve().parent.parent


def test_library_exports_every_declared_symbol():
    from flashinfer_b200 import _lib

    lib = _lib.load()
    header = (ROOT / "include" / "fi_b200.h").read_text()
    declared = set(re.findall(r"FI_API\s+[\w\s\*]+?\b(fi_\w+)\s*\(", header))
    assert declared, "no FI_API declarations found"
    for name in declared:
        assert hasattr(lib, name), f"libfi_b200.so does not export {name}"
    assert declared == set(_lib.PROTOTYPES), "ctypes prototypes out of sync with the header"
    assert lib.fi_version() >= 100
    assert lib.fi_plan_info_len() == 32


def lens_to_table(kv_lens, page_size):
    pages = [-(-l // page_size) for l in kv_lens]
    indptr = np.concatenate([[0], np.cumsum(pages)]).astype(np.int32)
    last = np.array([((l - 1) % page_size + 1) if l > 0 else 0 for l in kv_lens], dtype=np.int32)
    return indptr, last


CASES = [
    # (kv_lens, Hq, Hkv, page_size, num_ctas)
    ([4096] * 64, 32, 8, 16, 296),                       # C2
    ([8192] * 128, 8, 1, 16, 296),                       # C3 per rank
    ([131072] * 16, 8, 1, 16, 148),                      # C5 per rank
    ([1, 54, 97, 512, 2048, 0, 16384, 63, 64, 65], 32, 4, 8, 296),
    ([0, 9], 4, 1, 16, 148),                             # empty-request regression case
    ([0, 0, 0], 8, 2, 16, 148),
    ([7] * 17, 4, 4, 1, 148),
    ([5, 11, 130], 8, 8, 5, 31),                         # non-power-of-two page size
]


@pytest.mark.parametrize("head_group", ["1", "0"])
@pytest.mark.parametrize("with_indices", [False, True])
@pytest.mark.parametrize("kv_lens,hq,hkv,page,nctas", CASES)
def test_plan_visits_every_token_once(kv_lens, hq, hkv, page, nctas, head_group, with_indices, monkeypatch):
    monkeypatch.setenv("FI_B200_HEAD_GROUP", head_group)
    indptr, last = lens_to_table(kv_lens, page)
    # a shuffled page table handed to the planner: host_plan checks that every item head carries
    # exactly the page ids its first tiles gather and that the table is appended to the work list
    indices = np.random.default_rng(len(kv_lens)).permutation(int(indptr[-1]) + 3)[: int(indptr[-1])] \
        if with_indices else None
    rc, err, plan = host_plan(indptr, last, hq, hkv, 128, page, nctas, indices=indices)
    assert rc == 0, err
    replay_plan(plan, kv_lens, hkv)
    # work items: contiguous stream ranges of non-increasing size (guided self-scheduling);
    # CTA i starts on item i, the rest are drawn in order
    segs, iip = plan["segments"], plan["item_indptr"]
    nitems = len(iip) - 1
    loads = np.array([sum(int(s[3] - s[2]) for s in segs[iip[i]:iip[i + 1]]) for i in range(nitems)])
    total = sum((-(-l // TILE)) * hkv for l in kv_lens)
    assert loads.sum() == total
    assert plan["info"][PI["total_tiles"]] == total
    if total > 0:
        assert np.all(loads > 0)
        assert np.all(np.diff(loads[:-1]) <= 0), "item sizes must not grow along the stream"
        assert loads[-1] <= loads[max(nitems - 2, 0)]
        # equal-speed CTAs drawing items in order finish within a tail item of the ideal
        free = np.zeros(nctas, dtype=np.int64)
        for ld in loads:
            free[np.argmin(free)] += ld
        ideal = -(-total // nctas)
        assert free.max() <= max(ideal + max(int(loads[-2:].max()), 2), int(loads[0]))


def test_plan_head_groups_walk_the_same_request_ranges(monkeypatch):
    """Default stream order when Hkv divides the grid: the Hkv CTAs of a group hold the same
    (request, tile range) list, one KV head each — in an NHD cache they sweep whole pages together."""
    monkeypatch.setenv("FI_B200_HEAD_GROUP", "1")
    rng = np.random.default_rng(1)
    kv_lens = rng.integers(0, 9000, 40).tolist() + [0, 1, 64]
    hkv = 8
    indptr, last = lens_to_table(kv_lens, 16)
    rc, err, plan = host_plan(indptr, last, 32, hkv, 128, 16, 296)
    assert rc == 0, err
    replay_plan(plan, kv_lens, hkv)
    segs, iip = plan["segments"], plan["item_indptr"]
    nitems = len(iip) - 1
    assert nitems % hkv == 0
    loads = []
    for g in range(nitems // hkv):
        ref = [(int(s[0]), int(s[2]), int(s[3])) for s in segs[iip[g * hkv]:iip[g * hkv + 1]]]
        loads.append(sum(te - tb for _, tb, te in ref))
        for h in range(hkv):
            mine = segs[iip[g * hkv + h]:iip[g * hkv + h + 1]]
            assert all(int(s[1]) == h for s in mine)
            assert [(int(s[0]), int(s[2]), int(s[3])) for s in mine] == ref
    assert max(loads[:-1]) == min(loads[:-1]) and 0 < loads[-1] <= loads[0]
    # partials of a unit sit in consecutive slots in group (= chunk) order
    for r, h, sb, cnt in plan["merges"]:
        mine = sorted((int(s[2]), int(s[4])) for s in segs if int(s[0]) == r and int(s[1]) == h)
        assert [sl for _, sl in mine] == list(range(sb, sb + cnt))


def test_plan_uniform_batch_aligns_items_to_units():
    """bs 4 x kv 4096, 8 KV heads (64 tiles per unit, 2048 tiles, 296 CTAs): 7-tile items would straddle
    units; the planner takes 8-tile items instead — 256 items of exactly one segment each."""
    kv_lens = [4096] * 4
    indptr, last = lens_to_table(kv_lens, 16)
    rc, err, plan = host_plan(indptr, last, 32, 8, 128, 16, 296)
    assert rc == 0, err
    replay_plan(plan, kv_lens, 8)
    segs, iip = plan["segments"], plan["item_indptr"]
    assert len(iip) - 1 == 256
    assert all(iip[i + 1] - iip[i] == 1 for i in range(256))
    assert all(int(s[3] - s[2]) == 8 for s in segs)
    # many short requests: an item holds whole units (5 requests of 2 tiles -> 10-tile items, not 9 or 11)
    kv_lens = [128] * 1480
    indptr, last = lens_to_table(kv_lens, 16)
    rc, err, plan = host_plan(indptr, last, 8, 1, 128, 16, 296)
    assert rc == 0, err
    replay_plan(plan, kv_lens, 1)
    assert len(plan["merges"]) == 0


def test_plan_fp8_tiles_are_128_tokens():
    kv_lens = [8192] * 16 + [1, 127, 128, 129, 5000]
    indptr, last = lens_to_table(kv_lens, 16)
    rc, err, plan = host_plan(indptr, last, 8, 1, 128, 16, 296, q_dtype=1, kv_dtype=2)
    assert rc == 0, err
    replay_plan(plan, kv_lens, 1, tile=TILE_FP8)
    assert plan["info"][PI["total_tiles"]] == sum(-(-l // TILE_FP8) for l in kv_lens)


def test_plan_log_uniform_c4_balance():
    rng = np.random.default_rng(0)
    kv_lens = np.exp(rng.uniform(np.log(512), np.log(32768), 256)).round().astype(int).tolist()
    indptr, last = lens_to_table(kv_lens, 16)
    rc, err, plan = host_plan(indptr, last, 32, 8, 128, 16, 296)
    assert rc == 0, err
    replay_plan(plan, kv_lens, 8)
    assert len(plan["merges"]) <= 256 * 8


def test_plan_is_deterministic():
    kv_lens = [100, 2000, 31, 977]
    indptr, last = lens_to_table(kv_lens, 16)
    a = host_plan(indptr, last, 32, 8, 128, 16, 296)[2]
    b = host_plan(indptr, last, 32, 8, 128, 16, 296)[2]
    for key in ("segments", "merges", "item_indptr", "item_heads", "header"):
        assert np.array_equal(a[key], b[key])
    assert np.array_equal(a["info"], b["info"])


def test_plan_window_left():
    kv_lens = [300, 10, 1000]
    indptr, last = lens_to_table(kv_lens, 16)
    rc, err, plan = host_plan(indptr, last, 8, 2, 128, 16, 148, window_left=99)
    assert rc == 0, err
    replay_plan(plan, kv_lens, 2, window_left=99)


def test_plan_explicit_seq_lens():
    indptr = np.array([0, 4, 10], dtype=np.int32)
    rc, err, plan = host_plan(indptr, [16, 16], 8, 2, 128, 16, 148, kv_len=[50, 96])
    assert rc == 0, err
    replay_plan(plan, [50, 96], 2)


def test_plan_graph_mode_layout_is_shape_independent():
    """CUDA-graph contract: the launch grid and every offset that run() turns into a kernel
    parameter depend only on (bs, heads, num_ctas); the segment list (found by the kernel from
    the item heads) may move."""
    a = host_plan(*lens_to_table([100] * 8, 16), 32, 8, 128, 16, 296, cuda_graph=1)[2]["info"]
    b = host_plan(*lens_to_table([5000, 3, 77, 0, 12000, 640, 64, 1], 16), 32, 8, 128, 16, 296,
                  cuda_graph=1)[2]["info"]
    for key in ("num_ctas", "off_header", "off_item_heads", "off_work_counter", "off_counters",
                "off_merge", "off_tmp_o", "off_tmp_lse", "merge_grid", "stages"):
        assert a[PI[key]] == b[PI[key]], key


def test_plan_errors():
    indptr, last = lens_to_table([100], 16)
    rc, err, _ = host_plan(indptr, last, 30, 8, 128, 16, 148)
    assert rc == 1 and "multiple" in err
    # a GQA group of 16 is planned as two virtual KV heads of 8 query heads per real KV head
    rc, err, plan = host_plan(indptr, last, 32, 2, 128, 16, 148)
    assert rc == 0 and plan["info"][PI["num_kv_heads"]] == 4 and plan["info"][PI["kv_head_div"]] == 2
    replay_plan(plan, [100], 4)
    rc, err, _ = host_plan(indptr, last, 8, 2, 96, 16, 148)
    assert rc == 4 and "head_dim" in err
    rc, err, _ = host_plan(indptr, [17], 8, 2, 128, 16, 148)
    assert rc == 1 and "last_page_len" in err
    # split-KV partials do not fit the float workspace
    rc, err, _ = host_plan(*lens_to_table([100000], 16), 8, 2, 128, 16, 148, float_bytes=1024)
    assert rc == 2 and "workspace" in err.lower()
    rc, err, _ = host_plan(*lens_to_table([64] * 4096, 16), 8, 8, 128, 16, 148, int_bytes=4096)
    assert rc == 2 and "Increase the workspace buffer size" in err


def _unit_segments(plan, r, h):
    return [tuple(int(x) for x in s[2:7:1])[0:2] + (int(s[5]), int(s[6]))
            for s in plan["segments"] if s[0] == r and s[1] == h]


def test_plan_fixed_split_size_is_batch_invariant():
    """fixed_split_size (reference decode.py:880-885): a request's chunks — hence the arithmetic
    of its partials and the order they are merged in — do not depend on the rest of the batch."""
    page, split_pages = 16, 24                      # 384 tokens: not a multiple of the 64-token tile
    S = page * split_pages
    target = 5000
    alone = host_plan(*lens_to_table([target], page), 32, 8, 128, page, 296, fixed_split_pages=split_pages)[2]
    batch_lens = [77, 30000, target, 1, 0, 2048, 385, 384]
    batch = host_plan(*lens_to_table(batch_lens, page), 32, 8, 128, page, 296,
                      fixed_split_pages=split_pages)[2]
    replay_plan(alone, [target], 8, chunked=True)
    replay_plan(batch, batch_lens, 8, chunked=True)
    for h in range(8):
        a, b = _unit_segments(alone, 0, h), _unit_segments(batch, 2, h)
        assert a == b and len(a) == -(-target // S)
        for c, (tb, te, lo, hi) in enumerate(a):    # chunk c covers tokens [c S, (c+1) S)
            assert (lo, hi) == (c * S, min((c + 1) * S, target))
            assert (tb, te) == (lo // TILE, -(-hi // TILE))
    # one-chunk requests are written directly, multi-chunk requests always go through a merge
    merged = {(int(m[0]), int(m[1])): int(m[3]) for m in batch["merges"]}
    for r, L in enumerate(batch_lens):
        for h in range(8):
            chunks = -(-L // S) if L else 0
            assert merged.get((r, h), 1 if chunks else 0) == (chunks if chunks > 1 else (1 if chunks else 0))


def test_plan_disable_split_kv():
    kv_lens = [40000, 3, 0, 900, 64]
    plan = host_plan(*lens_to_table(kv_lens, 16), 16, 4, 128, 16, 296, disable_split_kv=1)[2]
    replay_plan(plan, kv_lens, 4)
    assert len(plan["merges"]) == 0 and plan["info"][PI["num_slots"]] == 0
    assert len(plan["segments"]) == len(kv_lens) * 4            # one whole-unit segment each
    assert all(int(s[4]) == -1 for s in plan["segments"])


def test_argument_block_entry_points_match_the_positional_ones():
    """fi_batch_decode_plan_args(FiDecodePlanArgs*) == fi_batch_decode_plan(...) bit for bit (host
    mode), and the ctypes mirrors have the C layout (sizes are checked against a compiled probe)."""
    import ctypes as C
    import subprocess
    import tempfile

    from flashinfer_b200 import _lib

    lib = _lib.load()
    kv_lens, hq, hkv, page, nctas = [4096] * 8 + [77, 0, 5000], 32, 8, 16, 296
    indptr, last = lens_to_table(kv_lens, page)
    idx = np.random.default_rng(1).permutation(int(indptr[-1])).astype(np.int32)
    rc, err, plan = host_plan(indptr, last, hq, hkv, 128, page, nctas, indices=idx)
    assert rc == 0, err
    ws = np.zeros(8 << 20, dtype=np.uint8)
    info = np.zeros(lib.fi_plan_info_len(), dtype=np.int64)
    a = _lib.DecodePlanArgs()
    a.float_ws, a.float_ws_bytes = None, 128 << 20
    a.int_ws = a.pinned_int_ws = ws.ctypes.data
    a.int_ws_bytes = ws.size
    a.indptr_host, a.last_page_len_host, a.kv_len_host = indptr.ctypes.data, last.ctypes.data, None
    a.batch_size, a.num_qo_heads, a.num_kv_heads, a.head_dim, a.page_size = len(kv_lens), hq, hkv, 128, page
    a.q_dtype = a.kv_dtype = 1
    a.window_left, a.cuda_graph, a.fixed_split_pages, a.disable_split_kv = -1, 0, 0, 0
    a.device, a.num_ctas_override, a.stream = -1, nctas, None
    a.kv_indices_host, a.num_indices = idx.ctypes.data, len(idx)
    a.plan_info_out = info.ctypes.data
    assert lib.fi_batch_decode_plan_args(C.byref(a)) == 0, _lib.last_error()
    assert np.array_equal(info, plan["info"])
    used = int(info[PI["int_bytes"]])
    rc2, _, plan2 = host_plan(indptr, last, hq, hkv, 128, page, nctas, indices=idx)
    assert used == plan2["info"][PI["int_bytes"]]
    assert lib.fi_batch_decode_plan_args(None) == _lib.FI_ERR_INVALID_ARGUMENT
    assert lib.fi_batch_decode_run_args(None) == _lib.FI_ERR_INVALID_ARGUMENT
    # struct layout: compile a three-line probe against the public header
    src = '#include <stdio.h>\n#include "fi_b200.h"\nint main(void){printf("%zu %zu\\n", sizeof(FiDecodePlanArgs), sizeof(FiDecodeRunArgs));return 0;}\n'
    with tempfile.TemporaryDirectory() as d:
        (Path(d) / "p.c").write_text(src)
        subprocess.run(["gcc", "-I", str(ROOT / "include"), "-o", f"{d}/p", f"{d}/p.c"], check=True)
        sizes = subprocess.run([f"{d}/p"], capture_output=True, text=True, check=True).stdout.split()
    assert [int(x) for x in sizes] == [C.sizeof(_lib.DecodePlanArgs), C.sizeof(_lib.DecodeRunArgs)]


@pytest.mark.parametrize("page", [8, 16, 64])
def test_head_pair_plan(page):
    """FI_PLAN_HEAD_PAIRS (include/fi_b200.h): fp8 KV at head_dim 128 is planned as (request, PAIR of KV
    heads) units over 64-token tiles — half the heads, every token of every pair visited exactly once —
    and every shape the flag does not apply to gets the plan it would have got without it."""
    from flashinfer_b200 import _lib

    kv_lens, hq, hkv, nctas = [4096] * 6 + [77, 0, 5000, 1, 64, 65], 64, 8, 296
    indptr, last = lens_to_table(kv_lens, page)
    idx = np.random.default_rng(3).permutation(int(indptr[-1])).astype(np.int32)
    rc, err, plan = host_plan(indptr, last, hq, hkv, 128, page, nctas, kv_dtype=_lib.DTYPE_E4M3, indices=idx,
                              flags=_lib.PLAN_HEAD_PAIRS)
    assert rc == 0, err
    info = plan["info"]
    assert info[PI["head_pairs"]] == 1 and info[PI["num_kv_heads"]] == hkv // 2 and info[PI["hq"]] == hq
    assert info[PI["total_tiles"]] == sum(-(-l // 64) for l in kv_lens) * (hkv // 2)
    replay_plan(plan, kv_lens, hkv // 2, tile=64)
    # not applicable: 16-bit KV, other head dims, an odd number of KV heads, groups above 8, tiny pages
    for kw in (dict(kv_dtype=1), dict(kv_dtype=_lib.DTYPE_E4M3, head_dim=256), dict(kv_dtype=_lib.DTYPE_E4M3, hkv=1),
               dict(kv_dtype=_lib.DTYPE_E4M3, hq=128, hkv=8)):
        hd = kw.get("head_dim", 128)
        args = (indptr, last, kw.get("hq", hq), kw.get("hkv", hkv), hd, page, nctas)
        rc1, err1, with_flag = host_plan(*args, kv_dtype=kw["kv_dtype"], indices=idx, flags=_lib.PLAN_HEAD_PAIRS)
        rc0, err0, without = host_plan(*args, kv_dtype=kw["kv_dtype"], indices=idx)
        assert rc1 == 0 and rc0 == 0, (err1, err0)
        assert with_flag["info"][PI["head_pairs"]] == 0
        assert np.array_equal(with_flag["info"], without["info"])
        assert np.array_equal(with_flag["segments"], without["segments"])
    indptr4, last4 = lens_to_table(kv_lens, 4)
    rc, err, small = host_plan(indptr4, last4, hq, hkv, 128, 4, nctas, kv_dtype=_lib.DTYPE_E4M3,
                               flags=_lib.PLAN_HEAD_PAIRS)
    assert rc == 0 and small["info"][PI["head_pairs"]] == 0


def test_wide_unit_plans():
    """FI_PLAN_WIDE_UNITS (include/fi_b200.h): up to 16 query columns per unit — GQA groups of 9 .. 16 are one
    unit per KV head instead of two virtual heads (16-bit KV at head_dim <= 128, fp8 KV at 128, TMA-sized pages);
    groups above 16 split into units of <= 16; every other shape gets the plan it would have got without the flag."""
    from flashinfer_b200 import _lib

    kv_lens, page, nctas = [4096] * 6 + [77, 0, 5000, 1], 16, 296
    indptr, last = lens_to_table(kv_lens, page)

    def plan(hq, hkv, d=128, kvd=1, flags=_lib.PLAN_WIDE_UNITS, pg=page, ip=indptr, lp=last):
        rc, err, pl = host_plan(ip, lp, hq, hkv, d, pg, nctas, kv_dtype=kvd, flags=flags)
        assert rc == 0, err
        return pl

    for hq, hkv, kvd, d in ((128, 8, 1, 128), (12, 1, 1, 64), (40, 4, 1, 128), (128, 8, _lib.DTYPE_E4M3, 128)):
        pl = plan(hq, hkv, d=d, kvd=kvd)
        info = pl["info"]
        assert info[PI["num_kv_heads"]] == hkv and info[PI["kv_head_div"]] == 1
        replay_plan(pl, kv_lens, hkv, tile=64 if kvd == 1 else 128)
    # 24 / 1 heads: two units of 12 columns; 64 / 1: four of 16
    assert plan(24, 1)["info"][PI["kv_head_div"]] == 2 and plan(64, 1)["info"][PI["kv_head_div"]] == 4
    # not applicable: head_dim 256, fp8 at head_dim 256, small pages -> as without the flag
    indptr4, last4 = lens_to_table(kv_lens, 4)
    for kw in (dict(d=256), dict(d=256, kvd=_lib.DTYPE_E4M3), dict(pg=4, ip=indptr4, lp=last4)):
        a, b = plan(128, 8, **kw), plan(128, 8, flags=0, **kw)
        assert a["info"][PI["kv_head_div"]] == 2
        assert np.array_equal(a["info"], b["info"]) and np.array_equal(a["segments"], b["segments"])
    # groups of <= 8 are untouched by the flag
    a, b = plan(32, 8), plan(32, 8, flags=0)
    assert np.array_equal(a["info"], b["info"]) and np.array_equal(a["segments"], b["segments"])


def _fuzz_plan_case(data):
    from hypothesis import strategies as st
    from flashinfer_b200 import _lib

    bs = data.draw(st.integers(1, 24), label="bs")
    len_st = st.one_of(st.sampled_from([0, 1, 63, 64, 65, 127, 128, 129, 4096]), st.integers(0, 9000))
    if data.draw(st.booleans(), label="uniform"):
        kv_lens = [data.draw(len_st, label="len")] * bs
    else:
        kv_lens = [data.draw(len_st, label="len") for _ in range(bs)]
    page = data.draw(st.sampled_from([1, 2, 4, 5, 8, 16, 32, 64, 128]), label="page")
    hkv = data.draw(st.sampled_from([1, 2, 3, 4, 8]), label="hkv")
    group = data.draw(st.sampled_from([1, 2, 3, 4, 5, 6, 7, 8, 9, 12, 16, 24, 32]), label="group")
    head_dim = data.draw(st.sampled_from([64, 128, 256, 512]), label="head_dim")
    fp8 = data.draw(st.booleans(), label="fp8")
    nctas = data.draw(st.sampled_from([1, 7, 31, 148, 296, 600]), label="nctas")
    window = data.draw(st.sampled_from([-1, -1, 0, 17, 1000]), label="window_left")
    mode = data.draw(st.sampled_from(["default", "default", "fixed", "nosplit"]), label="mode")
    flags = data.draw(st.sampled_from([0, _lib.PLAN_HEAD_PAIRS, _lib.PLAN_WIDE_UNITS,
                                       _lib.PLAN_HEAD_PAIRS | _lib.PLAN_WIDE_UNITS]), label="flags")
    graph = data.draw(st.booleans(), label="cuda_graph")
    with_indices = data.draw(st.booleans(), label="with_indices")
    return dict(kv_lens=kv_lens, page=page, hkv=hkv, hq=hkv * group, head_dim=head_dim,
                kv_dtype=_lib.DTYPE_E4M3 if fp8 else 1, nctas=nctas, window=window, mode=mode, flags=flags,
                graph=int(graph), with_indices=with_indices,
                fixed=data.draw(st.sampled_from([1, 3, 16]), label="fixed_split_pages") if mode == "fixed" else 0)


def _check_fuzzed_plan(c):
    from flashinfer_b200 import _lib

    indptr, last = lens_to_table(c["kv_lens"], c["page"])
    indices = None
    if c["with_indices"]:
        indices = np.random.default_rng(len(c["kv_lens"])).permutation(int(indptr[-1]) + 3)[: int(indptr[-1])]
    rc, err, plan = host_plan(indptr, last, c["hq"], c["hkv"], c["head_dim"], c["page"], c["nctas"],
                              kv_dtype=c["kv_dtype"], window_left=c["window"], cuda_graph=c["graph"],
                              fixed_split_pages=c["fixed"], disable_split_kv=int(c["mode"] == "nosplit"),
                              indices=indices, flags=c["flags"], int_bytes=32 << 20, float_bytes=1 << 36)
    assert rc == 0, err
    info = plan["info"]
    fp8 = c["kv_dtype"] == _lib.DTYPE_E4M3
    pairs = int(info[PI["head_pairs"]])
    div = int(info[PI["kv_head_div"]])
    tile = (TILE_FP8 if fp8 else TILE) // (2 if pairs else 1)
    unit_heads = int(info[PI["num_kv_heads"]])
    # what a unit is: a KV head, `div` virtual heads of one, or a pair of KV heads — never two of those at once
    assert unit_heads == (c["hkv"] // 2 if pairs else c["hkv"] * div) and not (pairs and div > 1)
    group = c["hq"] // c["hkv"]
    cols = group * 2 if pairs else group // div
    assert group % div == 0 and info[PI["hq"]] == c["hq"]
    wide = bool(c["flags"] & _lib.PLAN_WIDE_UNITS)
    assert cols <= (16 if wide or pairs else 8), (cols, div, pairs)
    if c["head_dim"] == 512:
        assert cols <= 4
    if pairs:
        assert fp8 and c["head_dim"] == 128 and c["flags"] & _lib.PLAN_HEAD_PAIRS and c["hkv"] % 2 == 0 \
            and group <= 8 and c["page"] >= 8 and c["page"] & (c["page"] - 1) == 0
    replay_plan(plan, c["kv_lens"], unit_heads, window_left=c["window"], tile=tile, chunked=c["mode"] == "fixed")
    if c["mode"] == "nosplit":
        assert len(plan["merges"]) == 0 and np.all(plan["segments"][:, 4] < 0)
    # a plan is a pure function of its arguments
    rc2, _, again = host_plan(indptr, last, c["hq"], c["hkv"], c["head_dim"], c["page"], c["nctas"],
                              kv_dtype=c["kv_dtype"], window_left=c["window"], cuda_graph=c["graph"],
                              fixed_split_pages=c["fixed"], disable_split_kv=int(c["mode"] == "nosplit"),
                              indices=indices, flags=c["flags"], int_bytes=32 << 20, float_bytes=1 << 36)
    assert rc2 == 0 and np.array_equal(again["info"], info) and np.array_equal(again["segments"], plan["segments"]) \
        and np.array_equal(again["item_heads"], plan["item_heads"]) and np.array_equal(again["merges"], plan["merges"])


def test_plan_fuzz():
    """Property test over the whole planner argument space (lengths incl. empty and tile-edge requests, page sizes,
    head shapes, GQA groups 1 .. 32, head dims, fp8 / 16-bit KV, grid sizes, sliding windows, fixed_split_size /
    disable_split_kv, FI_PLAN_* flags, CUDA-graph layout, with and without the host page table): every plan is
    replayed — each attended token of each unit exactly once, each unit written once, each partial slot owned by
    one merge, inline segments / page ids equal to the lists — and planned twice (bit-identical)."""
    from hypothesis import given, settings, HealthCheck, strategies as st

    @settings(max_examples=250, deadline=None, derandomize=True, suppress_health_check=list(HealthCheck))
    @given(st.data())
    def run(data):
        _check_fuzzed_plan(_fuzz_plan_case(data))

    run()
