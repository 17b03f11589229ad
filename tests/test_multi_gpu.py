"""GPU, >= 2 devices: the real multi-GPU path.  Every rank runs the CUDA kernels on its KV-head
shard of the SAME batch (views of one replicated cache), the output heads are gathered with the
product's collectives (NCCL all-gather, and the peer-write gather when it can be set up), and rank 0
compares the gathered o / lse with the CPU oracle on the unsharded inputs.  Skipped on a
one-GPU box; `gpurun --gpus 2 -- python -m pytest tests/test_multi_gpu.py -m gpu` runs it.
(tests/test_sharding.py covers the same host logic on gloo without a GPU.)"""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from helpers import make_paged_case, make_q
from oracle import attention_ref as orc

pytestmark = pytest.mark.gpu


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, layout, kv_dtype, ret):
    import flashinfer_b200 as fi
    from flashinfer_b200 import sharding

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        kv_lens = [97, 1, 4099, 0, 2048, 130, 16384, 777]
        hq, hkv, d, page = 32, 8, 128, 16
        q_dtype = torch.bfloat16
        cache32, indptr, indices, last = make_paged_case(kv_lens, hkv, d, page, layout, torch.float32, seed=3)
        is8 = kv_dtype == torch.float8_e4m3fn
        cache = cache32.to(kv_dtype)
        k_scale, v_scale = (0.5, 0.25) if is8 else (None, None)
        q = make_q(len(kv_lens), hq, d, q_dtype, seed=4)
        sh = sharding.head_shard(hq, hkv, rank, world)
        # a production cache is allocated per rank with Hkv / N heads: make the shard contiguous
        cache_loc = sharding.shard_paged_kv_cache(cache, layout, sh).contiguous().to(dev)
        q_loc = sharding.shard_q(q, sh).to(dev)          # strided view semantics are covered elsewhere
        ws = torch.zeros(64 << 20, dtype=torch.uint8, device=dev)
        w = fi.BatchDecodeWithPagedKVCacheWrapper(ws, layout)
        w.plan(indptr, indices, last, sh.num_qo_heads, sh.num_kv_heads, d, page, q_data_type=q_dtype,
               kv_data_type=kv_dtype)
        o_loc, lse_loc = w.run(q_loc, cache_loc, k_scale=k_scale, v_scale=v_scale, return_lse=True)
        o_all = sharding.gather_heads(o_loc, world)
        lse_all = sharding.gather_heads(lse_loc, world)
        peer = None
        if hasattr(sharding, "PeerGather"):
            pg = sharding.PeerGather((len(kv_lens), hq, d), q_dtype, dev, sh)
            for _ in range(3):  # both buffers of the ping-pong and a wrap-around
                peer = pg.gather(o_loc)
                torch.cuda.synchronize()
                assert torch.equal(peer, o_all), "peer-write gather differs from the NCCL all-gather"
        torch.cuda.synchronize()
        if rank == 0:
            o_ref, lse_ref = orc.batch_decode_ref(q, cache, indptr, indices, last, layout,
                                                  k_scale=k_scale or 1.0, v_scale=v_scale or 1.0)
            err_o = (o_all.float().cpu() - o_ref).abs().max().item()
            err_l = (lse_all.cpu() - lse_ref).abs().max().item()
            assert err_o <= (3e-2 if is8 else 1e-2), err_o
            assert err_l <= 1e-3, err_l
        dist.barrier()
        ret[rank] = "ok" if peer is not None or not hasattr(sharding, "PeerGather") else "no peer gather"
    except BaseException as e:  # noqa: BLE001
        ret[rank] = f"{type(e).__name__}: {e}"
        raise
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("layout,kv_dtype", [("NHD", torch.bfloat16), ("HND", torch.float8_e4m3fn)])
def test_head_sharded_decode_on_real_gpus(layout, kv_dtype):
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    world = 8 if n >= 8 else (4 if n >= 4 else 2)
    port = _free_port()
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, port, layout, kv_dtype, ret), nprocs=world, join=True)
    assert dict(ret) == {r: "ok" for r in range(world)}
