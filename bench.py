#!/usr/bin/env python
"""bench.py — paged-KV batch decode benchmark (driver contract).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--workload c2|c3|c4|c5] [--page-size P] [--no-extras] [--no-cpu]

A "step" is one `BatchDecodeWithPagedKVCacheWrapper.run` (decode kernel, split-KV merge fused)
over one synthetic batch; the metric is BASELINE.json's "paged decode us/step" (plus tokens/s and
HBM TB/s in the same JSON line).  At N GPUs the paged cache is sharded by KV head (north_star):
every rank decodes Hkv/N KV heads of every request, no collective inside the timed region; the
gather of O across head shards is timed and reported separately (`allgather_us`: NCCL
all-gather; `peer_gather_us`: the peer-write gather of flashinfer_b200/sharding.py).

Blocks of the JSON line beyond the driver contract:
  parity          the (gathered) o / lse of sampled requests against the CPU oracle on the unsharded
                  inputs, after the timed legs; the process exits non-zero when it fails
  other_workloads C3 / C4 / C5 (and, at N=1, the per-rank shapes of an 8-way sharding), same method
  comparators     N=1: the image's upstream flashinfer-python (library code, never on the product
                  path) on the same tensors: fa2 (CUDA-core decode), fa2_tc (use_tensor_cores=True),
                  trtllm-gen, cuDNN — each {us, tbps, loaded, error}
  sample_config   the reference's published decode sample (benchmarks/samples/
                  sample_testlist_output.csv:14-23: 64/8 heads, D128, page 16, bs 1 / 16, random
                  kv_len <= 1024, CUDA-graph timed), ours beside the comparators and the CSV numbers

`--impl reference`: the reference has no CPU implementation of this path (it is a CUDA library).
When oracle/_ref/liboracle_ref.so is present (the reference's own DecodePlan + decode + merge
templates compiled unmodified from its headers, see oracle/Makefile) and a GPU is visible, the
reference arm times THAT on the same workload and the same per-rank shape: its plan + its kernels
with its default launch attributes (PDL on), nothing of ours on the path.  Otherwise (or with
--cpu-reference) it times the oracle port (oracle/attention_ref.py = the reference test-suite's
fp32 PyTorch attention) on the host cores over a bounded sample of the workload, scaled to the
full batch; that CPU figure is also what `cpu_baseline` reports.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import sys
import threading
import time
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

PAGE = 16  # BASELINE's page size (--page-size changes it for the page-geometry probes)
HEAD_DIM = 128


# --------------------------------------------------------------------------- workloads
def workload(name: str):
    """BASELINE.json configs (C2..C5).  Returns dict(kv_lens, hq, hkv, kv_dtype)."""
    if name == "c2":
        return dict(desc="Llama-3-8B GQA 32/8 d128 bs64 kv4096 page16 bf16 NHD", kv_lens=[4096] * 64,
                    hq=32, hkv=8, kv_dtype="bf16")
    if name == "c3":
        return dict(desc="Llama-3-70B 64/8 d128 FP8-e4m3 KV bs128 kv8192 page16 NHD", kv_lens=[8192] * 128,
                    hq=64, hkv=8, kv_dtype="e4m3")
    if name == "c4":
        rng = np.random.default_rng(0)
        lens = np.exp(rng.uniform(np.log(512), np.log(32768), 256)).round().astype(int).tolist()
        return dict(desc="bs256 kv_len log-uniform[512,32768] 32/8 d128 page16 bf16 NHD", kv_lens=lens,
                    hq=32, hkv=8, kv_dtype="bf16")
    if name == "c5":
        return dict(desc="Llama-3-70B 64/8 d128 bs16 kv131072 page16 bf16 NHD", kv_lens=[131072] * 16,
                    hq=64, hkv=8, kv_dtype="bf16")
    raise SystemExit(f"unknown workload {name}")


def algorithmic_bytes(kv_lens, hq, hkv, kv_es, d=HEAD_DIM) -> int:
    """SURVEY.md §8(d) / reference flashinfer/testing/utils.py:518-531: K+V bytes at the actual
    lengths + q + o (page-table ints, lse and split-KV partials are overhead, not counted)."""
    return sum(kv_lens) * hkv * d * 2 * kv_es + len(kv_lens) * hq * d * (2 + 2)


def page_table(kv_lens, page=PAGE, seed=0, shuffle=True):
    pages = [-(-l // page) for l in kv_lens]
    total = sum(pages)
    g = torch.Generator().manual_seed(seed)
    # shuffled: exercises the gather
    indices = (torch.randperm(total, generator=g) if shuffle else torch.arange(total)).to(torch.int32)
    indptr = torch.zeros(len(kv_lens) + 1, dtype=torch.int32)
    indptr[1:] = torch.tensor(np.cumsum(pages), dtype=torch.int32)
    last = torch.tensor([((l - 1) % page + 1) if l > 0 else 0 for l in kv_lens], dtype=torch.int32)
    return indptr, indices, last, total


# --------------------------------------------------------------------------- clocks sampler
class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML from before the warm-up to the end of
    the timed region; `mark()` tags the start of the timed region so both parts are reported."""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop_evt = threading.Event()
        self._mark = None
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:  # noqa: BLE001
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {
            nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
            nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
            nv.nvmlClocksThrottleReasonHwPowerBrakeSlowdown: "hw_power_brake",
        }
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, nm in names.items():
                    if r & bit:
                        self.reasons.add(nm)
            except Exception:  # noqa: BLE001
                pass
            time.sleep(0.0005)

    def mark(self):
        self._mark = len(self.samples)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=1)
        m = self._mark if self._mark is not None else 0
        timed, warm = self.samples[m:], self.samples[:m]
        # a timed region shorter than one NVML query still has the sample that straddles it
        if not timed and self.samples:
            timed = self.samples[-1:]
        return {"sm_mhz": float(np.median(timed)) if timed else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(timed),
                "sm_mhz_warmup": float(np.median(warm)) if warm else None, "samples_warmup": len(warm),
                "sampled": "NVML every ~0.5 ms; `samples` = timed region (+ the isolated-launch leg of our arm), "
                           "`samples_warmup` = set-up and warm-up before it"}


# --------------------------------------------------------------------------- one workload on one rank
class Case:
    """Synthetic inputs of one workload shard, resident in HBM, behind our public wrapper.
    Distinct cache copies are rotated when one copy would fit in the 126 MB L2 several times over,
    so every step streams from HBM."""

    def __init__(self, name, wl, hq, hkv, dev, kv_layout="NHD", page=16, seed=1234, shuffle=True):
        import flashinfer_b200 as fi

        self.name, self.wl, self.hq, self.hkv, self.dev, self.page = name, wl, hq, hkv, dev, page
        self.kv_layout = kv_layout
        self.kv_lens = wl["kv_lens"]
        self.bs = len(self.kv_lens)
        self.kv_dtype = torch.float8_e4m3fn if wl["kv_dtype"] == "e4m3" else torch.bfloat16
        self.kv_es = 1 if wl["kv_dtype"] == "e4m3" else 2
        self.q_dtype = torch.bfloat16
        indptr, indices, last, self.total_pages = page_table(self.kv_lens, page, shuffle=shuffle)
        self.indptr, self.indices, self.last = indptr.pin_memory(), indices.pin_memory(), last.pin_memory()
        self.cache_bytes = self.total_pages * 2 * page * hkv * HEAD_DIM * self.kv_es
        self.n_copies = max(1, min(8, math.ceil((512 << 20) / self.cache_bytes)))
        g = torch.Generator(device=dev).manual_seed(seed)
        self.caches = []
        for _ in range(self.n_copies):
            shape = (self.total_pages, 2, page, hkv, HEAD_DIM) if kv_layout == "NHD" else \
                (self.total_pages, 2, hkv, page, HEAD_DIM)
            c = torch.randn(*shape, device=dev, dtype=torch.bfloat16, generator=g)
            self.caches.append(c.to(self.kv_dtype) if self.kv_es == 1 else c)
            del c
        # per-tensor scales (values are synthetic)
        self.k_scale, self.v_scale = (0.5, 0.5) if self.kv_es == 1 else (None, None)
        gc = torch.Generator().manual_seed(seed)
        self.q_host = torch.randn(self.bs, hq, HEAD_DIM, generator=gc).to(self.q_dtype).pin_memory()
        self.o_host = torch.empty(self.bs, hq, HEAD_DIM, dtype=self.q_dtype).pin_memory()
        self.q = self.q_host.to(dev)
        self.out = torch.empty_like(self.q)
        self.ws = torch.zeros(128 << 20, dtype=torch.uint8, device=dev)
        self.wrapper = fi.BatchDecodeWithPagedKVCacheWrapper(self.ws, kv_layout)

    def plan(self):
        self.wrapper.plan(self.indptr, self.indices, self.last, self.hq, self.hkv, HEAD_DIM, self.page,
                          q_data_type=self.q_dtype, kv_data_type=self.kv_dtype)

    def step(self, i):
        self.wrapper.run(self.q, self.caches[i % self.n_copies], k_scale=self.k_scale,
                         v_scale=self.v_scale, out=self.out)

    @property
    def bytes_rank(self):
        return algorithmic_bytes(self.kv_lens, self.hq, self.hkv, self.kv_es)

    def l2_note(self):
        return (f"KV per rank {self.cache_bytes / 2**20:.0f} MiB > L2; {self.n_copies} cache "
                f"cop{'y' if self.n_copies == 1 else 'ies'} rotated")


def time_steps(step, steps, warmup, sync_all, sampler=None):
    """W untimed warm-up steps (topped up to 30: the first launches after an idle GPU run at ramping
    clocks), then K steps between two events on the launching stream."""
    for i in range(max(warmup, 30)):
        step(i)
    sync_all()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if sampler is not None:
        sampler.mark()
    e0.record()
    for i in range(steps):
        step(i)
    e1.record()
    sync_all()
    return e0.elapsed_time(e1) * 1e3 / steps


def isolated_launch_us(step, n, dev):
    """The same kernel timed one launch at a time behind an L2 flush that READS a 256 MB buffer
    (the L2 then holds someone else's clean lines, as between a model's GEMMs)."""
    flush = torch.zeros(64 << 20, dtype=torch.float32, device=dev)  # 256 MB, twice the L2
    per = []
    for i in range(n):
        flush.sum()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        step(i)
        b.record()
        b.synchronize()
        per.append(a.elapsed_time(b))
    return float(np.median(per)) * 1e3


def e2e_us(case, n, host_out=False, fused=False):
    """End to end through the public API with HOST buffers: q H2D from pinned memory + plan (page
    table from pinned host memory) + run + o D2H + sync, every step.  host_out: run(out = the pinned
    host tensor) — the kernel's epilogue writes o over PCIe, no separate D2H copy.  fused: the wrapper's
    step() — the same plan and run through ONE library call (fi_batch_decode_step), host out."""
    ts = []
    if fused:  # step() re-uses the scales of the last run() and the cache tensor bound by it
        case.plan()
        case.wrapper.run(case.q, case.caches[0], k_scale=case.k_scale, v_scale=case.v_scale, out=case.o_host)
    for i in range(n + 2):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        case.q.copy_(case.q_host, non_blocking=True)  # enqueued first: the copy runs while the host plans
        if fused:
            case.wrapper.step(case.indptr, case.indices, case.last, case.q, case.caches[0], out=case.o_host)
            torch.cuda.synchronize()
            ts.append(time.perf_counter() - t0)
            continue
        case.plan()
        if host_out:
            case.wrapper.run(case.q, case.caches[i % case.n_copies], k_scale=case.k_scale, v_scale=case.v_scale,
                             out=case.o_host)
        else:
            case.step(i)
            case.o_host.copy_(case.out, non_blocking=True)
        torch.cuda.synchronize()
        ts.append(time.perf_counter() - t0)
    return float(np.median(ts[2:])) * 1e6


# --------------------------------------------------------------------------- parity against the oracle
def parity_check(case, world, rank, requests):
    """o / lse of `requests` from a fresh run() — gathered across the head shards with the product's
    own gather when world > 1 — against oracle.batch_decode_ref (CPU fp32) on the UNSHARDED inputs
    (q and the requests' KV pages collected from every rank).  Returns the JSON block on rank 0."""
    import torch.distributed as dist

    from flashinfer_b200 import sharding

    reqs = sorted(set(int(r) for r in requests))
    o, lse = case.wrapper.run(case.q, case.caches[0], k_scale=case.k_scale, v_scale=case.v_scale,
                              return_lse=True)
    ip = case.indptr.tolist()
    pages = torch.cat([case.indices[ip[r]:ip[r + 1]] for r in reqs]).long().to(case.dev)
    kv = case.caches[0][pages]  # [n, 2, page, hkv, D] (NHD) or [n, 2, hkv, page, D] (HND)
    if case.kv_layout == "HND":
        kv = kv.transpose(2, 3)
    kv = kv.contiguous()
    sel = torch.tensor(reqs, device=case.dev)
    q_s, o_s, lse_s = case.q[sel].contiguous(), o[sel].contiguous(), lse[sel].contiguous()
    if world > 1:
        o_s = sharding.gather_heads(o_s, world)        # the collective under test
        lse_s = sharding.gather_heads(lse_s, world)
        q_s = sharding.gather_heads(q_s, world)
        kv8 = kv.view(torch.uint8) if case.kv_es == 1 else kv
        parts = [torch.empty_like(kv8) for _ in range(world)]
        dist.all_gather(parts, kv8)
        kv = torch.cat(parts, dim=3)
        if case.kv_es == 1:
            kv = kv.view(case.kv_dtype)
    torch.cuda.synchronize()
    if rank != 0:
        return None
    from oracle import attention_ref as orc

    n_pages = [ip[r + 1] - ip[r] for r in reqs]
    indptr2 = [0] + list(np.cumsum(n_pages))
    last2 = [int(case.last[r]) for r in reqs]
    hkv_all, hq_all = kv.shape[3], q_s.shape[1]
    heads = list(range(hkv_all))
    if sum(case.kv_lens[r] for r in reqs) * hq_all > 6_000_000:
        # very long requests (C5): the fp32 oracle on every head would take minutes of host time;
        # check the first and the last KV head (rank 0's and rank N-1's shard) with all their query heads
        heads = sorted({0, hkv_all - 1})
        grp = hq_all // hkv_all
        qh = [h * grp + j for h in heads for j in range(grp)]
        kv, q_s, o_s, lse_s = kv[:, :, :, heads], q_s[:, qh], o_s[:, qh], lse_s[:, qh]
    o_ref, lse_ref = orc.batch_decode_ref(q_s.cpu(), kv.cpu(), indptr2, list(range(sum(n_pages))), last2, "NHD",
                                          k_scale=case.k_scale or 1.0, v_scale=case.v_scale or 1.0)
    err_o = (o_s.float().cpu() - o_ref).abs().max().item()
    err_l = (lse_s.cpu() - lse_ref).abs().max().item()
    tol_o = 3e-2 if case.kv_es == 1 else 1e-2
    return {"ok": bool(err_o <= tol_o and err_l <= 1e-3), "max_abs_o": err_o, "max_abs_lse": err_l,
            "tol_o": tol_o, "tol_lse": 1e-3, "requests": reqs, "kv_heads": heads,
            "checked": "run() output" + (f", all-gathered over {world} head shards," if world > 1 else "") +
                       " vs oracle.batch_decode_ref (CPU fp32) on the unsharded q / KV of those requests"}


# --------------------------------------------------------------------------- CPU baseline / reference arm
def cpu_reference_time(wl, sample_requests: int, reps: int, page=16):
    """Time the oracle port on the host cores on `sample_requests` requests of the workload
    (full KV length, all heads) and scale to the whole batch.  Returns (us_per_full_step, info)."""
    from oracle import attention_ref as orc

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    kv_lens = wl["kv_lens"][:sample_requests]
    hq, hkv = wl["hq"], wl["hkv"]
    indptr, indices, last, total = page_table(kv_lens, page, seed=1)
    g = torch.Generator().manual_seed(0)
    dt = torch.bfloat16
    cache = torch.randn(total, 2, page, hkv, HEAD_DIM, generator=g).to(dt)
    if wl["kv_dtype"] == "e4m3":
        cache = cache.to(torch.float8_e4m3fn)
    q = torch.randn(len(kv_lens), hq, HEAD_DIM, generator=g).to(dt)
    orc.batch_decode_ref(q, cache, indptr, indices, last, "NHD", requests=[0])  # warm-up
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        orc.batch_decode_ref(q, cache, indptr, indices, last, "NHD")
        ts.append(time.perf_counter() - t0)
    scale = sum(wl["kv_lens"]) / max(sum(kv_lens), 1)
    us = float(np.median(ts)) * 1e6 * scale
    info = {"cores": cores, "kind": "port",
            "sample": f"{len(kv_lens)} of {len(wl['kv_lens'])} requests at full kv_len, all heads, "
                      f"median of {reps}, scaled x{scale:.1f} by KV tokens"}
    return us, info


def config_block(args, wl, world, hq, hkv, cache_bytes, n_copies):
    """Both arms print this same `config` (same keys, same values for the same run)."""
    return {"workload": f"{args.workload}: {wl['desc']}" +
                        (" [cache laid out HND for this run]" if args.kv_layout != "NHD" else "") +
                        (f" [page_size {args.page_size}]" if args.page_size != 16 else ""),
            "sharding": f"kv-head x{world}" if args.shard_of <= 1 else
                        f"ONE rank of a kv-head x{args.shard_of} sharding (not the whole job)",
            "heads_per_rank": [hq, hkv], "page_table": args.page_order,
            "l2": f"KV per rank {cache_bytes / 2**20:.0f} MiB > L2; {n_copies} cache "
                  f"cop{'y' if n_copies == 1 else 'ies'} rotated"}


def run_reference_gpu(args, wl, so_path):
    """The reference's own plan + CUDA kernels (oracle/_ref) on one GPU.  At N > 1 the workload is
    sharded by KV head exactly like our arm and rank 0's shard is timed: every rank runs the same
    shape and there is no collective in the timed region, so rank 0's time is the job's."""
    import ctypes as C

    from flashinfer_b200 import sharding

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    ref = C.CDLL(str(so_path))
    ref.ref_last_error.restype = C.c_char_p
    world = max(args.gpus, 1)
    page = args.page_size
    shard = sharding.head_shard(wl["hq"], wl["hkv"], 0, world if args.shard_of <= 1 else args.shard_of)
    kv_lens, hq, hkv = wl["kv_lens"], shard.num_qo_heads, shard.num_kv_heads
    bs = len(kv_lens)
    kv_dtype = torch.float8_e4m3fn if wl["kv_dtype"] == "e4m3" else torch.bfloat16
    kv_es = 1 if wl["kv_dtype"] == "e4m3" else 2
    indptr, indices, last, total_pages = page_table(kv_lens, page, shuffle=args.page_order == "shuffled")
    cache_bytes = total_pages * 2 * page * hkv * HEAD_DIM * kv_es
    n_copies = max(1, min(8, math.ceil((512 << 20) / cache_bytes)))
    torch.manual_seed(1234)
    caches = []
    for _ in range(n_copies):
        c = torch.randn(total_pages, 2, page, hkv, HEAD_DIM, device=dev, dtype=torch.bfloat16)
        caches.append(c.to(kv_dtype) if kv_es == 1 else c)
        del c
    q_host = torch.randn(bs, hq, HEAD_DIM).to(torch.bfloat16).pin_memory()
    o_host = torch.empty(bs, hq, HEAD_DIM, dtype=torch.bfloat16).pin_memory()
    q = q_host.to(dev)
    out = torch.empty_like(q)
    fws = torch.zeros(256 << 20, dtype=torch.uint8, device=dev)
    iws = torch.zeros(8 << 20, dtype=torch.uint8, device=dev)
    pws = torch.zeros(8 << 20, dtype=torch.uint8).pin_memory()
    tbl_host = [t.pin_memory() for t in (indptr, indices, last)]
    tbl_dev = [t.to(dev) for t in tbl_host]
    info = torch.zeros(16, dtype=torch.int64)
    vp = C.c_void_p
    kvd = 2 if kv_es == 1 else 1
    stream = torch.cuda.current_stream().cuda_stream
    pdl = int(ref.ref_get_pdl()) if hasattr(ref, "ref_get_pdl") else 0

    def call(mode, i=0):
        k, v = caches[i % n_copies][:, 0], caches[i % n_copies][:, 1]
        rc = ref.ref_batch_decode_mode(
            vp(fws.data_ptr()), C.c_size_t(fws.numel()), vp(iws.data_ptr()), vp(pws.data_ptr()),
            C.c_size_t(iws.numel()), vp(indptr.data_ptr()), bs, hq, hkv, HEAD_DIM, page, 1, kvd, 0, 0,
            vp(q.data_ptr()), C.c_int64(q.stride(0)), C.c_int64(q.stride(1)), vp(k.data_ptr()),
            vp(v.data_ptr()), C.c_int64(k.stride(0)), C.c_int64(k.stride(1)), C.c_int64(k.stride(2)),
            vp(tbl_dev[1].data_ptr()), vp(tbl_dev[0].data_ptr()), vp(tbl_dev[2].data_ptr()),
            vp(out.data_ptr()), None, C.c_float(1.0 / math.sqrt(HEAD_DIM)), vp(stream),
            vp(info.data_ptr()), mode)
        if rc != 0:
            raise RuntimeError(f"reference shim failed: {ref.ref_last_error().decode()}")

    call(1)
    sampler = ClockSampler(0)
    sampler.start()
    us_step = time_steps(lambda i: call(2, i), args.steps, args.warmup, torch.cuda.synchronize, sampler)
    clocks = sampler.stop()
    # end to end like our arm: page table + q from pinned host memory, plan + run, o back to the host
    t_e2e = []
    for i in range(max(3, min(args.steps, 20)) + 2):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for d_t, h_t in zip(tbl_dev, tbl_host):
            d_t.copy_(h_t, non_blocking=True)
        call(1)
        q.copy_(q_host, non_blocking=True)
        call(2, i)
        o_host.copy_(out, non_blocking=True)
        torch.cuda.synchronize()
        t_e2e.append(time.perf_counter() - t0)
    e2e = float(np.median(t_e2e[2:])) * 1e6
    h2d = q_host.numel() * 2 + sum(t.numel() for t in tbl_host) * 4
    cpu_us, cpu_info = cpu_reference_time(wl, sample_requests=1, reps=1, page=page)
    line = {
        "impl": "reference", "metric": "paged_decode_us_per_step", "value": us_step, "unit": "us",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "warmup_run": max(args.warmup, 30),
        "ms_per_step": us_step / 1e3,
        "higher_is_better": False, "scaling": "strong", "vs_baseline": None,
        "dtype": "bf16" if kv_es == 2 else "bf16 q/o, e4m3 kv", "data": "synthetic",
        "tokens_per_s": bs / (us_step * 1e-6),
        "config": config_block(args, wl, world, hq, hkv, cache_bytes, n_copies),
        "reference_impl": {
            "what": "the reference's own DecodePlan + BatchDecodeWithPagedKVCache (+ merge) templates (its "
                    "CUDA-core `fa2` decode, the kernel behind use_tensor_cores=False), compiled unmodified "
                    "from its headers into oracle/_ref; the reference has no CPU implementation of this path",
            "enable_pdl": bool(pdl), "launches_per_step": int(info[10]),
            "sharded_like_ours": "rank 0's KV-head shard on one GPU; all ranks run the same shape and the "
                                 "timed region has no collective" if world > 1 else "unsharded"},
        "cpu_baseline": {"value": cpu_us, "unit": "us", **cpu_info},
        "e2e": {"value": e2e, "unit": "us", "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": o_host.numel() * 2,
                "what": "page table H2D + reference plan + q H2D + reference kernels + o D2H"},
        "gpu_launches": int(info[10]) * args.steps, "clocks": clocks,
    }
    print(json.dumps(line))


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    wl = workload(args.workload)
    so = ROOT / "oracle" / "_ref" / "liboracle_ref.so"
    if not args.cpu_reference and so.exists() and torch.cuda.is_available():
        return run_reference_gpu(args, wl, so)
    us, info = cpu_reference_time(wl, sample_requests=2, reps=max(1, min(args.steps, 30)), page=args.page_size)
    bs = len(wl["kv_lens"])
    line = {
        "impl": "reference", "metric": "paged_decode_us_per_step", "value": us, "unit": "us",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": us / 1e3,
        "higher_is_better": False, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "tokens_per_s": bs / (us * 1e-6),
        "config": {"workload": f"{args.workload}: {wl['desc']}"},
        "cpu_baseline": {"value": us, "unit": "us", **info},
        "e2e": {"value": us, "unit": "us", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


# --------------------------------------------------------------------------- upstream comparators (N = 1)
def _upstream():
    """The image's upstream flashinfer-python, with its JIT pointed at the modules pre-built by
    scripts/build_upstream_comparators.py (baseline/_ref/fi_aot) so nothing compiles on the box."""
    os.environ.setdefault("FLASHINFER_WORKSPACE_BASE", "/tmp/fi_b200_upstream_ws")
    import flashinfer as up
    from flashinfer.jit import env as jit_env

    aot = ROOT / "baseline" / "_ref" / "fi_aot"
    if not aot.exists():
        raise RuntimeError("baseline/_ref/fi_aot is missing (run scripts/build_upstream_comparators.py where "
                           "nvcc is; a JIT build inside the bench would take minutes)")
    jit_env.FLASHINFER_AOT_DIR = aot
    from flashinfer.jit import core as jit_core

    def _no_jit(self, *a, **k):  # a module that was not pre-built must not start a compile here
        raise RuntimeError(f"upstream module {self.name} is not pre-built in baseline/_ref/fi_aot")

    jit_core.JitSpec.build = _no_jit
    return up


def _graph_time_us(fn, iters_in_graph=10, replays=30):
    """The reference harness's method (flashinfer/testing/utils.py bench_gpu_time_with_cudagraph):
    `iters_in_graph` back-to-back calls captured in one CUDA graph, median replay time / iters."""
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(3):
            fn()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(iters_in_graph):
            fn()
    for _ in range(5):
        g.replay()
    torch.cuda.synchronize()
    ts = []
    for _ in range(replays):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        g.replay()
        b.record()
        b.synchronize()
        ts.append(a.elapsed_time(b) * 1e3 / iters_in_graph)
    return float(np.median(ts))


def upstream_runners(up, dev, hq, hkv, page, indptr, indices, last, kv_lens, q, cache_nhd, kv_dtype,
                     k_scale=None, v_scale=None, which=("fa2", "fa2_tc", "trtllm-gen", "cudnn")):
    """{name: (callable -> out tensor) | Exception} for the upstream decode backends on our tensors.
    fa2 / fa2_tc read the NHD cache in place; trtllm-gen and cuDNN get an HND copy (their layout)."""
    runners = {}
    bs = len(kv_lens)
    indptr_d, indices_d, last_d = indptr.to(dev), indices.to(dev), last.to(dev)
    cache_hnd = None

    def hnd():
        nonlocal cache_hnd
        if cache_hnd is None:
            cache_hnd = cache_nhd.transpose(2, 3).contiguous()
        return cache_hnd

    for name in which:
        try:
            if name in ("fa2", "fa2_tc", "trtllm-gen"):
                layout = "HND" if name == "trtllm-gen" else "NHD"
                ws = torch.zeros(128 << 20, dtype=torch.uint8, device=dev)
                w = up.BatchDecodeWithPagedKVCacheWrapper(
                    ws, layout, use_tensor_cores=(name != "fa2"),
                    backend=name if name == "trtllm-gen" else "auto")
                w.plan(indptr_d if name != "trtllm-gen" else indptr_d.clone(), indices_d, last_d, hq, hkv,
                       HEAD_DIM, page, q_data_type=q.dtype, kv_data_type=kv_dtype)
                cache = hnd() if layout == "HND" else cache_nhd
                o = torch.empty_like(q)
                runners[name] = (lambda w=w, cache=cache, o=o: w.run(q, cache, k_scale=k_scale,
                                                                     v_scale=v_scale, out=o))
            elif name == "cudnn":
                if kv_dtype != q.dtype:
                    raise RuntimeError("cuDNN decode comparator is set up for 16-bit KV only")
                c = hnd()
                k_c, v_c = c[:, 0], c[:, 1]
                max_pages = max(-(-l // page) for l in kv_lens)
                bt = torch.zeros(bs, max_pages, dtype=torch.int32)
                ip = indptr.tolist()
                for i in range(bs):
                    bt[i, :ip[i + 1] - ip[i]] = indices[ip[i]:ip[i + 1]]
                bt = bt.to(dev)
                lens = torch.tensor(kv_lens, dtype=torch.int32, device=dev).view(bs, 1, 1, 1)
                ws = torch.empty(128 << 20, dtype=torch.int8, device=dev)
                ragged = (torch.arange(0, bs + 1, device=dev) * (hq * HEAD_DIM)).long()
                o = torch.empty_like(q)
                runners[name] = (lambda k_c=k_c, v_c=v_c, ws=ws, lens=lens, bt=bt, ragged=ragged, o=o:
                                 up.decode.cudnn_batch_decode_with_kv_cache(
                                     q, k_c, v_c, 1.0 / math.sqrt(HEAD_DIM), ws, max_sequence_kv=max(kv_lens),
                                     actual_seq_lens_kv=lens, block_tables=bt, is_cuda_graph_compatible=True,
                                     batch_offsets_q=ragged, batch_offsets_o=ragged, out=o))
            r = runners[name]()
            torch.cuda.synchronize()
            del r
        except Exception as e:  # noqa: BLE001
            runners[name] = e
    return runners


def comparators_block(case):
    """Upstream flashinfer-python backends on the main workload's tensors (cache copy 0), each timed
    like `value` (K back-to-back run() calls between events after a warm-up); never on the product path."""
    out = {"library": None, "note": "image's upstream flashinfer-python (library code), same q / page table / "
                                    "cache as our arm; trtllm-gen and cuDNN read an HND copy of the cache"}
    try:
        up = _upstream()
        out["library"] = f"flashinfer-python {up.__version__}"
    except Exception as e:  # noqa: BLE001
        out["error"] = f"import failed: {type(e).__name__}: {e}"[:300]
        return out
    case.plan()
    ours = case.wrapper.run(case.q, case.caches[0], k_scale=case.k_scale, v_scale=case.v_scale).float()
    runners = upstream_runners(up, case.dev, case.hq, case.hkv, case.page, case.indptr, case.indices, case.last,
                               case.kv_lens, case.q, case.caches[0], case.kv_dtype, case.k_scale, case.v_scale)
    for name, r in runners.items():
        if isinstance(r, Exception):
            out[name] = {"loaded": False, "error": f"{type(r).__name__}: {r}"[:300]}
            continue
        try:
            o = r()
            torch.cuda.synchronize()
            diff = (o.float() - ours).abs().max().item()
            us = time_steps(lambda i: r(), 50, 10, torch.cuda.synchronize)
            out[name] = {"loaded": True, "us": us, "tbps": case.bytes_rank / (us * 1e-6) / 1e12,
                         "max_abs_vs_ours": diff}
        except Exception as e:  # noqa: BLE001
            out[name] = {"loaded": True, "error": f"{type(e).__name__}: {e}"[:300]}
    return out


def sample_config_block(dev):
    """The reference's published decode sample (benchmarks/samples/sample_testlist.txt:11-12,
    sample_testlist_output.csv:14-23): Llama-3.1-70B head shape 64/8, D128, page 16, bf16, bs 1 and 16,
    kv_len random <= 1024, timed by CUDA-graph replay of 10 back-to-back calls (its harness's method)."""
    import flashinfer_b200 as fi

    published = {1: {"fa2": 34.82, "fa2_tc": 10.45, "cudnn": 10.68, "trtllm-gen": 5.73,
                     "hardware": "B200 (reference's sample run)"},
                 16: {"fa2": 55.30, "fa2_tc": 17.21, "cudnn": 13.76, "trtllm-gen": 9.63,
                      "hardware": "B200 (reference's sample run)"}}
    res = {"method": "CUDA graph of 10 back-to-back run() calls, median of 30 replays / 10 (warm L2, as the "
                     "reference's bench_gpu_time_with_cudagraph)", "cases": []}
    try:
        up = _upstream()
    except Exception:  # noqa: BLE001
        up = None
    hq, hkv, page = 64, 8, 16
    for bs in (1, 16):
        rng = np.random.default_rng(42 + bs)
        kv_lens = rng.integers(1, 1025, bs).tolist()
        indptr, indices, last, total = page_table(kv_lens, page, shuffle=False)
        g = torch.Generator(device=dev).manual_seed(42)
        cache = torch.randn(total, 2, page, hkv, HEAD_DIM, device=dev, dtype=torch.bfloat16, generator=g)
        q = torch.randn(bs, hq, HEAD_DIM, device=dev, dtype=torch.bfloat16, generator=g)
        ws = torch.zeros(128 << 20, dtype=torch.uint8, device=dev)
        w = fi.BatchDecodeWithPagedKVCacheWrapper(ws, "NHD")
        w.plan(indptr, indices, last, hq, hkv, HEAD_DIM, page, q_data_type=torch.bfloat16)
        o = torch.empty_like(q)
        entry = {"batch_size": bs, "avg_kv_len": float(np.mean(kv_lens)),
                 "ours_us": _graph_time_us(lambda: w.run(q, cache, out=o)), "published_us": published[bs]}
        ours = o.float().clone()
        if up is not None:
            runners = upstream_runners(up, dev, hq, hkv, page, indptr, indices, last, kv_lens, q, cache,
                                       torch.bfloat16)
            for name, r in runners.items():
                if isinstance(r, Exception):
                    entry[name] = {"loaded": False, "error": f"{type(r).__name__}: {r}"[:200]}
                    continue
                try:
                    d = (r().float() - ours).abs().max().item()
                    entry[name] = {"loaded": True, "us": _graph_time_us(r), "max_abs_vs_ours": d}
                except Exception as e:  # noqa: BLE001
                    entry[name] = {"loaded": True, "error": f"{type(e).__name__}: {e}"[:200]}
        res["cases"].append(entry)
    return res


# --------------------------------------------------------------------------- GPU arm
def measure_extra(name, wl, hq, hkv, dev, world, rank, sync_all, steps, shard_note):
    """One more workload (sharded like the main one), device-resident timing + parity on request 0."""
    case = Case(name, wl, hq, hkv, dev, seed=4321 + rank)
    case.plan()
    us = time_steps(case.step, steps, 10, sync_all)
    par = parity_check(case, world, rank, [0])
    vals = torch.tensor([us], device=dev, dtype=torch.float64)
    if world > 1:
        import torch.distributed as dist

        dist.all_reduce(vals, op=dist.ReduceOp.MAX)
    us = float(vals[0])
    peak, _ = hbm_peak_gbs()
    achieved = case.bytes_rank / (us * 1e-6) / 1e9
    block = {"workload": f"{name}: {wl['desc']}", "sharding": shard_note, "heads_per_rank": [hq, hkv],
             "value": us, "unit": "us", "steps": steps,
             "roofline": {"achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                          "frac_of_8tbps": achieved / 8000.0,
                          "algorithmic_bytes_per_launch": case.bytes_rank},
             "parity": par}
    del case
    torch.cuda.empty_cache()
    return block


def run_ours(args):
    import torch.distributed as dist

    from flashinfer_b200 import _lib, sharding

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch with torch.distributed.run for --gpus > 1")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    wl = workload(args.workload)
    hq_all, hkv_all = wl["hq"], wl["hkv"]
    try:
        # --shard-of N (development aid): time rank 0's slice of an N-way sharding on this one GPU
        shard = sharding.head_shard(hq_all, hkv_all, rank, world) if args.shard_of <= 1 else \
            sharding.head_shard(hq_all, hkv_all, 0, args.shard_of)
    except ValueError as e:
        raise SystemExit(str(e))
    hq, hkv = shard.num_qo_heads, shard.num_kv_heads
    lib = _lib.load()

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local_rank)
    sampler.start()  # before the warm-up; mark() tags the timed region
    if args.loader != "auto":  # A/B of the two loaders on the same page geometry (development aid)
        lib.fi_set_tuning(0, 0, 1 if args.loader == "tma" else 2, 0)
    case = Case(args.workload, wl, hq, hkv, dev, kv_layout=args.kv_layout, page=args.page_size, seed=1234 + rank,
                shuffle=args.page_order == "shuffled")
    bs = case.bs
    case.plan()

    # ---- device-resident timing: K steps between two events on the launching stream
    for i in range(max(args.warmup, 30)):
        case.step(i)
    sync_all()
    n0 = lib.fi_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sync_all()
    sampler.mark()
    e0.record()
    for i in range(args.steps):
        case.step(i)
    e1.record()
    sync_all()
    launches = lib.fi_launch_count() - n0
    ms_total = e0.elapsed_time(e1)

    kern_us = isolated_launch_us(case.step, min(args.steps, 50), dev)
    clocks = sampler.stop()  # timed region + the isolated-launch leg right after it (same kernel, same load)
    sync_all()
    # e2e: run(out = pinned host tensor) — the epilogue writes o over PCIe while the kernel is still
    # streaming; the copy-back form (run into HBM, then o D2H) is measured beside it
    e2e_copy_back = e2e_us(case, max(3, min(args.steps, 20)))
    e2e_plan_run = e2e_us(case, max(3, min(args.steps, 20)), host_out=True)
    torch.cuda.synchronize()
    if not torch.equal(case.o_host, case.out.cpu()):
        raise RuntimeError("run(out = pinned host tensor) differs from run(out = device tensor)")
    e2e = e2e_us(case, max(3, min(args.steps, 20)), fused=True)
    torch.cuda.synchronize()
    case.wrapper.run(case.q, case.caches[0], k_scale=case.k_scale, v_scale=case.v_scale, out=case.out)
    torch.cuda.synchronize()
    if not torch.equal(case.o_host, case.out.cpu()):
        raise RuntimeError("step() differs from plan() + run()")
    case.e2e_copy_back, case.e2e_plan_run = e2e_copy_back, e2e_plan_run
    h2d = case.q_host.numel() * 2 + (case.indptr.numel() + case.indices.numel() + case.last.numel()) * 4
    d2h = case.o_host.numel() * 2

    # ---- the only exchange: gather of O across head shards, timed separately (not part of `value`)
    ag_us = pg_us = None
    gather_info = None
    if world > 1:
        def timed(fn, n=20):
            for _ in range(5):
                fn()
            sync_all()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(n):
                fn()
            b.record()
            sync_all()
            return a.elapsed_time(b) / n * 1e3

        ag_us = timed(lambda: sharding.gather_heads(case.out, world))
        if hasattr(sharding, "PeerGather"):
            try:
                pg = sharding.PeerGather((bs, hq_all, HEAD_DIM), case.q_dtype, dev, shard)
                full_nccl = sharding.gather_heads(case.out, world)
                full_peer = pg.gather(case.out)
                torch.cuda.synchronize()
                same = bool(torch.equal(full_nccl, full_peer))
                pg_us = timed(lambda: pg.gather(case.out))
                gather_info = {"peer_equals_nccl": same, "transport": pg.transport}
                if not same:
                    raise SystemExit("peer-write gather differs from the NCCL all-gather")
            except SystemExit:
                raise
            except Exception as e:  # noqa: BLE001
                gather_info = {"peer_error": f"{type(e).__name__}: {e}"[:300]}

    # ---- parity of the (gathered) result against the CPU oracle on the unsharded inputs
    parity = parity_check(case, world, rank, [0, 1, bs // 2, bs - 1])

    # ---- max over ranks
    vals = torch.tensor([ms_total, kern_us, e2e, ag_us or 0.0, pg_us or 0.0], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(vals, op=dist.ReduceOp.MAX)
    ms_total, kern_us, e2e, ag_max, pg_max = [float(x) for x in vals.tolist()]

    extras = {}
    plain = args.page_size == 16 and args.page_order == "shuffled" and args.loader == "auto"
    if not args.no_extras and args.shard_of <= 1 and args.workload == "c2" and plain:
        steps_x = 30
        for nm in ("c3", "c4", "c5"):
            w2 = workload(nm)
            sh = sharding.head_shard(w2["hq"], w2["hkv"], rank, world)
            extras[nm] = measure_extra(nm, w2, sh.num_qo_heads, sh.num_kv_heads, dev, world, rank, sync_all,
                                       steps_x, f"kv-head x{world}")
        if world == 1:  # the per-rank shapes of the 8-GPU form of the configs, on this one GPU
            for nm in ("c2", "c3", "c5"):
                w2 = workload(nm)
                sh = sharding.head_shard(w2["hq"], w2["hkv"], 0, 8)
                extras[f"{nm}_rank_of_8"] = measure_extra(
                    nm, w2, sh.num_qo_heads, sh.num_kv_heads, dev, 1, 0, sync_all, steps_x,
                    "ONE rank of a kv-head x8 sharding, timed on one GPU")

    if rank == 0:
        line = emit(args, wl, world, ms_total, kern_us, e2e, h2d, d2h, launches, clocks, case)
        if world > 1:
            line["allgather_us"] = ag_max
            if pg_us is not None:
                line["peer_gather_us"] = pg_max
            if gather_info:
                line["gather"] = gather_info
        line["parity"] = parity
        if extras:
            line["other_workloads"] = extras
        if world == 1 and not args.no_extras and args.shard_of <= 1 and plain:
            line["comparators"] = comparators_block(case)
            if args.workload == "c2":
                line["sample_config"] = sample_config_block(dev)
        if world == 1 and not args.no_cpu:
            # a quarter of the C2 batch (~0.2 s per pass on 16 cores); 2 requests of the heavier workloads
            us, info = cpu_reference_time(wl, sample_requests=16 if args.workload == "c2" else 2, reps=3,
                                          page=args.page_size)
            line["cpu_baseline"] = {"value": us, "unit": "us", **info}
        print(json.dumps(line))
        bad = [k for k, v in [("main", parity)] + [(k, v.get("parity")) for k, v in extras.items()]
               if v is not None and not v["ok"]]
    else:
        bad = []
    if world > 1:
        flag = torch.tensor([len(bad)], device=dev)
        dist.broadcast(flag, 0)
        bad = bad or (["rank0"] if int(flag) else [])
        dist.destroy_process_group()
    if bad:
        raise SystemExit(f"parity check failed: {bad}")


def hbm_peak_gbs():
    """Roofline denominator: the driver-measured copy bandwidth, else the recipe's fallback."""
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:  # noqa: BLE001
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic_per_launch(workload_name, world):
    """dram bytes (read+write) per launch of the decode kernel from the committed `ncu --set full`
    summary of THIS workload on one GPU; null for any other configuration."""
    p = ROOT / "profiles" / "decode_kernel_traffic.json"
    if world != 1 or not p.exists():
        return None
    try:
        return json.loads(p.read_text()).get(workload_name, {}).get("dram_bytes_per_launch")
    except Exception:  # noqa: BLE001
        return None


def emit(args, wl, world, ms_total, kern_us, e2e, h2d, d2h, launches, clocks, case):
    bs = len(wl["kv_lens"])
    us_step = ms_total * 1e3 / args.steps
    bytes_rank = case.bytes_rank  # one launch on one rank
    peak, peak_src = hbm_peak_gbs()
    # one launch per step and nothing else in the timed region: the kernel's average launch
    # duration is the step time (events on the launching stream around the K steps)
    achieved = bytes_rank / (us_step * 1e-6) / 1e9
    traffic = ncu_traffic_per_launch(args.workload, world) \
        if args.page_size == 16 and args.kv_layout == "NHD" and args.shard_of <= 1 else None
    return {
        "metric": "paged_decode_us_per_step", "value": us_step, "unit": "us", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "warmup_run": max(args.warmup, 30),
        "ms_per_step": us_step / 1e3,
        "higher_is_better": False, "scaling": "strong", "vs_baseline": None,
        "dtype": "bf16" if case.kv_es == 2 else "bf16 q/o, e4m3 kv", "data": "synthetic",
        "tokens_per_s": bs / (us_step * 1e-6),
        "hbm_tbps_whole_job": bytes_rank * world / (us_step * 1e-6) / 1e12,
        "config": config_block(args, wl, world, case.hq, case.hkv, case.cache_bytes, case.n_copies),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "frac_of_8tbps": achieved / 8000.0,
                     "traffic": traffic, "peak_source": peak_src,
                     "kernel": "paged_decode_kernel / paged_decode_tc5_kernel (split-KV merge fused), "
                               "one launch per run()",
                     "kernel_us": us_step, "kernel_us_isolated": kern_us,
                     "algorithmic_bytes_per_launch": bytes_rank},
        "e2e": {"value": e2e, "unit": "us", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "plan_run_us": getattr(case, "e2e_plan_run", None),
                "copy_back_us": getattr(case, "e2e_copy_back", None),
                "what": "q H2D (pinned, enqueued first so it overlaps the host planner) + wrapper.step(page table in "
                        "pinned host memory, q, cache, out = pinned host tensor) + sync, per step: step() is plan "
                        "(work list + indices H2D) and run in one library call, the kernel epilogue stores o over "
                        "PCIe (checked bit-equal to plan() + run() into HBM); plan_run_us = the same through separate "
                        "plan() and run(out = pinned host) calls; copy_back_us = plan() + run() into HBM + o D2H copy"},
        "gpu_launches": int(launches), "clocks": clocks,
    }


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=["c2", "c3", "c4", "c5"])
    ap.add_argument("--page-size", type=int, default=16,
                    help="page size of the synthetic cache (BASELINE's configs use 16; 1 is SGLang's default)")
    ap.add_argument("--page-order", default="shuffled", choices=["shuffled", "sequential"],
                    help="page table of the synthetic cache: a random permutation (default) or the identity")
    ap.add_argument("--loader", default="auto", choices=["auto", "tma", "cpasync"],
                    help="development aid: force one of the two KV loaders of the 16-bit kernel")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-extras", action="store_true",
                    help="skip other_workloads / comparators / sample_config (development runs, ncu)")
    ap.add_argument("--kv-layout", default="NHD", choices=["NHD", "HND"],
                    help="paged cache layout (BASELINE's configs are NHD; HND is the fast one for fp8 KV)")
    ap.add_argument("--shard-of", type=int, default=1,
                    help="development aid: run only rank 0's slice of an N-way KV-head sharding on one GPU")
    ap.add_argument("--cpu-reference", action="store_true",
                    help="--impl reference: time the CPU oracle port even when oracle/_ref and a GPU exist")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
