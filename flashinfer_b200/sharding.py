"""KV-head sharding of paged decode across GPUs (one process per GPU).

Decode attention is independent per KV head: rank r of N owns KV heads
[r Hkv/N, (r+1) Hkv/N) of every page and the query heads of their GQA groups, plans and runs
its slice with the ordinary single-GPU wrapper, and the only exchange is the all-gather of the
output heads (what a tensor-parallel layer feeds to its row-parallel o_proj anyway, so serving
stacks usually skip even that).  The reference has no multi-GPU decode path of its own
(SURVEY.md §8e): frameworks on top of it shard exactly like this.

Everything up to `gather_heads` is host-side index arithmetic plus one torch.distributed collective;
it runs unchanged on the gloo backend with CPU tensors, which is how tests/test_sharding.py covers it.
`PeerGather` is the NVLink-native form of that collective: one kernel of libfi_b200.so
(csrc/gather.cu), chained to the decode launch, that stores the rank's heads straight into every
rank's [bs, Hq, D] buffer (NVLS multicast or peer-mapped stores) and returns when all ranks' heads
have arrived — no NCCL call, no transpose copy, no allocation per call.
"""
from __future__ import annotations

import ctypes as C
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


class PeerGather:
    """Peer-write gather of the output heads: `gather(o_local)` returns this rank's `[bs, Hq, D]`
    view holding every rank's heads, in rank (= head) order, exactly what `gather_heads` returns.

    Set-up (once): two symmetric `[bs, Hq, D]` buffers (ping-pong by call parity) and a 256-byte
    control block per rank, allocated with torch's symmetric memory (CUDA VMM handles exchanged over
    the process group's store) so that every rank has every rank's buffers mapped, plus the NVLS
    multicast address when the fabric offers one.  Per call: ONE launch of `fi_peer_gather` on the
    current stream (programmatic dependent launch after the decode kernel), no host synchronisation.

    Why two buffers are enough: a rank starts call t + 2 (which overwrites the buffer of call t) only
    after its call t + 1 completed, and that needed every rank's flag for t + 1, which a rank raises
    from a kernel that follows — in its stream — whatever consumed the result of call t.  Callers
    that read the returned view from ANOTHER stream must order that read before their next-but-one
    `gather` themselves.  The returned tensor is a view of library-owned memory: copy it if it must
    outlive the next-but-one call."""

    def __init__(self, full_shape, dtype: torch.dtype, device: torch.device, shard: HeadShard,
                 group: Optional[dist.ProcessGroup] = None, use_multicast: Optional[bool] = None):
        import os

        import torch.distributed._symmetric_memory as symm_mem

        from . import _lib

        if not torch.device(device).type == "cuda":
            raise ValueError("PeerGather needs CUDA tensors (use gather_heads on the CPU)")
        if shard.world > 16:
            raise ValueError("PeerGather supports at most 16 ranks")
        bs, hq, d = full_shape
        if hq != shard.num_qo_heads * shard.world:
            raise ValueError("full_shape[1] must be world * heads per rank")
        self.shard, self.dtype, self.device = shard, dtype, torch.device(device)
        self.full_shape = (bs, hq, d)
        self._lib, self._check = _lib.load(), _lib.check
        grp = group if group is not None else dist.group.WORLD
        self._buf = symm_mem.empty((2, bs, hq, d), dtype=dtype, device=self.device)
        self._ctl = symm_mem.empty((64,), dtype=torch.int32, device=self.device)
        self._ctl.zero_()
        torch.cuda.synchronize(self.device)
        self._hbuf = symm_mem.rendezvous(self._buf, grp)
        self._hctl = symm_mem.rendezvous(self._ctl, grp)
        self._hctl.barrier()  # every rank's control block is zero before anybody's first flag
        world = shard.world
        half = bs * hq * d * self._buf.element_size()
        self._dst = [(C.c_void_p * world)(*[int(p) + k * half for p in self._hbuf.buffer_ptrs]) for k in (0, 1)]
        self._ctl_ptrs = (C.c_void_p * world)(*[int(p) for p in self._hctl.buffer_ptrs])
        if use_multicast is None:
            use_multicast = os.environ.get("FI_B200_GATHER_MULTICAST", "1") != "0"
        mc = int(getattr(self._hbuf, "multicast_ptr", 0) or 0) if use_multicast else 0
        self._mc = [mc + k * half if mc else None for k in (0, 1)]
        self.transport = "nvls multicast (multimem.st)" if mc else "peer-mapped stores (P2P)"
        self._row_bytes = shard.num_qo_heads * d * self._buf.element_size()
        self._stride = hq * d * self._buf.element_size()
        self._epoch = 0
        self._dev_idx = self.device.index if self.device.index is not None else torch.cuda.current_device()

    def gather(self, o_local: torch.Tensor, enable_pdl: bool = True) -> torch.Tensor:
        from .utils import current_stream_ptr

        bs, hq, d = self.full_shape
        if tuple(o_local.shape) != (bs, self.shard.num_qo_heads, d) or o_local.dtype != self.dtype:
            raise ValueError(f"o_local must be [{bs}, {self.shard.num_qo_heads}, {d}] {self.dtype}")
        if not o_local.is_contiguous():
            raise ValueError("o_local must be contiguous")
        self._epoch += 1
        k = self._epoch & 1
        self._check(self._lib.fi_peer_gather(
            o_local.data_ptr(), self._dst[k], self._mc[k], self._ctl_ptrs, self.shard.world, self.shard.rank,
            bs, self._row_bytes, self._stride, self.shard.rank * self._row_bytes,
            self._epoch & 0xFFFFFFFF, int(bool(enable_pdl)), self._dev_idx, current_stream_ptr(self.device)))
        return self._buf[k]

    def status(self) -> int:
        """0, or 1 + the rank whose flag a wait gave up on after 2 s (synchronises the device)."""
        return int(self._ctl[33].item())
