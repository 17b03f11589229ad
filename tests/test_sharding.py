"""KV-head sharding across ranks (flashinfer_b200/sharding.py), world_size 2 on the gloo backend
with CPU tensors: every rank plans its head slice through the C-ABI planner (host mode), computes
its slice with the CPU oracle standing in for the kernel, and the all-gathered output must equal
the unsharded oracle bit for bit."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from helpers import host_plan, make_paged_case, make_q, replay_plan
from oracle import attention_ref as orc


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, layout, tuple_cache, q_dtype, ret):
    from flashinfer_b200 import sharding

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.set_num_threads(2)
        kv_lens = [97, 1, 512, 0, 2048, 130]
        hq, hkv, d, page = 16, 4, 128, 16
        cache, indptr, indices, last = make_paged_case(kv_lens, hkv, d, page, layout, q_dtype, seed=3)
        q = make_q(len(kv_lens), hq, d, q_dtype, seed=4)
        sh = sharding.head_shard(hq, hkv, rank, world)
        assert (sh.num_qo_heads, sh.num_kv_heads) == (hq // world, hkv // world)
        # --- the rank's plan covers its own slice: every (request, local head) token once
        rc, err, plan = host_plan(indptr.numpy(), last.numpy(), sh.num_qo_heads, sh.num_kv_heads, d,
                                  page, 296)
        assert rc == 0, err
        replay_plan(plan, kv_lens, sh.num_kv_heads)
        assert plan["info"][13] == sh.num_kv_heads
        # --- the rank's slice of q and of the cache are views, not copies
        q_loc = sharding.shard_q(q, sh)
        pkv = (cache[:, 0], cache[:, 1]) if tuple_cache else cache
        pkv_loc = sharding.shard_paged_kv_cache(pkv, layout, sh)
        first = pkv_loc[0] if tuple_cache else pkv_loc
        assert first.untyped_storage().data_ptr() == cache.untyped_storage().data_ptr()
        cache_loc = torch.stack(pkv_loc, dim=1) if tuple_cache else pkv_loc
        # --- compute (oracle in place of the kernel), gather, compare with the unsharded result
        o_loc, lse_loc = orc.batch_decode_ref(q_loc, cache_loc, indptr, indices, last, layout)
        o_all = sharding.gather_heads(o_loc.to(q_dtype), world)
        lse_all = sharding.gather_heads(lse_loc, world)
        o_ref, lse_ref = orc.batch_decode_ref(q, cache, indptr, indices, last, layout)
        assert o_all.shape == (len(kv_lens), hq, d) and lse_all.shape == (len(kv_lens), hq)
        assert torch.equal(o_all, o_ref.to(q_dtype)), (o_all.float() - o_ref).abs().max()
        assert torch.equal(lse_all, lse_ref)
        # out= variant and world == 1 passthrough
        buf = torch.empty_like(o_all)
        assert sharding.gather_heads(o_loc.to(q_dtype), world, out=buf) is buf and torch.equal(buf, o_all)
        assert sharding.gather_heads(o_loc, 1) is o_loc
        ret[rank] = "ok"
    except BaseException as e:  # noqa: BLE001
        ret[rank] = f"{type(e).__name__}: {e}"
        raise
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("layout,tuple_cache,q_dtype", [
    ("NHD", False, torch.float16), ("HND", False, torch.bfloat16), ("NHD", True, torch.bfloat16)])
def test_head_sharded_decode_world2_gloo(layout, tuple_cache, q_dtype):
    world = 2
    port = _free_port()
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, port, layout, tuple_cache, q_dtype, ret), nprocs=world, join=True)
    assert dict(ret) == {0: "ok", 1: "ok"}


def test_head_shard_arithmetic_and_errors():
    from flashinfer_b200 import sharding

    seen_q, seen_kv = [], []
    for r in range(4):
        sh = sharding.head_shard(64, 8, r, 4)
        assert (sh.num_qo_heads, sh.num_kv_heads) == (16, 2)
        seen_q += list(range(64))[sh.qo_slice]
        seen_kv += list(range(8))[sh.kv_slice]
        # a rank's query heads are exactly the GQA groups of its KV heads
        assert [h // 8 for h in range(64)[sh.qo_slice]] == sorted(list(range(8))[sh.kv_slice] * 8)
    assert seen_q == list(range(64)) and seen_kv == list(range(8))
    with pytest.raises(ValueError):
        sharding.head_shard(32, 8, 0, 3)
    with pytest.raises(ValueError):
        sharding.head_shard(30, 8, 0, 2)
    with pytest.raises(ValueError):
        sharding.head_shard(32, 8, 2, 2)
    with pytest.raises(KeyError):
        sharding.shard_paged_kv_cache(torch.zeros(1, 2, 4, 2, 8), "XYZ", sharding.head_shard(4, 2, 0, 2))
    c = torch.arange(2 * 2 * 4 * 4 * 8).reshape(2, 2, 4, 4, 8)
    assert torch.equal(sharding.shard_paged_kv_cache(c, "NHD", sharding.head_shard(8, 4, 1, 2)), c[:, :, :, 2:4])
    assert torch.equal(sharding.shard_paged_kv_cache(c, "HND", sharding.head_shard(8, 4, 1, 2)), c[:, :, 2:4])
