"""KV-head sharding of paged decode across GPUs (one process per GPU).

Decode attention is independent per KV head: rank r of N owns KV heads
[r Hkv/N, (r+1) Hkv/N) of every page and the query heads of their GQA groups, plans and runs
its slice with the ordinary single-GPU wrapper, and the only exchange is the all-gather of the
output heads (what a tensor-parallel layer feeds to its row-parallel o_proj anyway, so serving
stacks usually skip even that).  The reference has no multi-GPU decode path of its own
(SURVEY.md §8e): frameworks on top of it shard exactly like this.

Everything here is host-side index arithmetic plus one torch.distributed collective; it runs
unchanged on the gloo backend with CPU tensors, which is how tests/test_sharding.py covers it.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Tuple, Union

import torch
import torch.distributed as dist


@dataclass(frozen=True)
class HeadShard:
    rank: int
    world: int
    num_qo_heads: int   # local
    num_kv_heads: int   # local
    qo_begin: int       # first global query head of this rank
    kv_begin: int       # first global KV head of this rank

    @property
    def qo_slice(self) -> slice:
        return slice(self.qo_begin, self.qo_begin + self.num_qo_heads)

    @property
    def kv_slice(self) -> slice:
        return slice(self.kv_begin, self.kv_begin + self.num_kv_heads)


def head_shard(num_qo_heads: int, num_kv_heads: int, rank: int, world: int) -> HeadShard:
    """Contiguous KV-head ranges; the GQA groups of a rank's KV heads stay with it."""
    if world <= 0 or not 0 <= rank < world:
        raise ValueError(f"bad rank {rank} / world {world}")
    if num_qo_heads % num_kv_heads != 0:
        raise ValueError(f"num_qo_heads {num_qo_heads} is not a multiple of num_kv_heads {num_kv_heads}")
    if num_kv_heads % world != 0:
        raise ValueError(f"num_kv_heads {num_kv_heads} is not divisible by {world} ranks")
    group = num_qo_heads // num_kv_heads
    hkv = num_kv_heads // world
    return HeadShard(rank, world, hkv * group, hkv, rank * hkv * group, rank * hkv)


def shard_q(q: torch.Tensor, shard: HeadShard) -> torch.Tensor:
    """q [bs, Hq, D] -> this rank's heads (a strided view: run() takes q strides)."""
    return q[:, shard.qo_slice]


def shard_paged_kv_cache(
    paged_kv_cache: Union[torch.Tensor, Tuple[torch.Tensor, torch.Tensor]],
    kv_layout: str,
    shard: HeadShard,
):
    """This rank's KV heads of a paged cache, as views (no copy): the page table is shared by all
    ranks, only the head axis is cut.  A production cache is ALLOCATED per rank with Hkv/N heads;
    this is for tests and for carving a replicated cache."""
    if kv_layout not in ("NHD", "HND"):
        raise KeyError(f"Invalid kv_layout {kv_layout}")
    if isinstance(paged_kv_cache, tuple):
        k, v = paged_kv_cache
        ax = 2 if kv_layout == "NHD" else 1
        return (k.narrow(ax, shard.kv_begin, shard.num_kv_heads),
                v.narrow(ax, shard.kv_begin, shard.num_kv_heads))
    ax = 3 if kv_layout == "NHD" else 2
    return paged_kv_cache.narrow(ax, shard.kv_begin, shard.num_kv_heads)


def gather_heads(o_local: torch.Tensor, world: int, group: Optional[dist.ProcessGroup] = None,
                 out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """All-gather of the per-rank output heads: [bs, Hq/N, ...] -> [bs, Hq, ...] in rank order.
    One collective (NCCL over NVLink for CUDA tensors, gloo for CPU tensors)."""
    if world == 1:
        return o_local
    o_local = o_local.contiguous()
    stacked = torch.empty((world,) + tuple(o_local.shape), dtype=o_local.dtype, device=o_local.device)
    if o_local.is_cuda:
        dist.all_gather_into_tensor(stacked, o_local, group=group)
    else:  # gloo has no all_gather_into_tensor for every dtype: use the list form
        parts = [stacked[i] for i in range(world)]
        if o_local.dtype in (torch.bfloat16, torch.float16):
            parts32 = [torch.empty(o_local.shape, dtype=torch.float32) for _ in range(world)]
            dist.all_gather(parts32, o_local.float(), group=group)
            for dst, src in zip(parts, parts32):
                dst.copy_(src)
        else:
            dist.all_gather(parts, o_local, group=group)
    # [N, bs, h, ...] -> [bs, N, h, ...] -> [bs, N h, ...]
    res = stacked.movedim(0, 1).reshape((o_local.shape[0], world * o_local.shape[1]) + tuple(o_local.shape[2:]))
    if out is not None:
        out.copy_(res)
        return out
    return res
