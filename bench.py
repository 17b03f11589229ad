#!/usr/bin/env python
"""bench.py — paged-KV batch decode benchmark (driver contract).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2|c3|c4|c5]

A "step" is one `BatchDecodeWithPagedKVCacheWrapper.run` (decode kernel + split-KV merge) over
one synthetic batch; the metric is BASELINE.json's "paged decode us/step" (plus tokens/s and
HBM TB/s in the same JSON line).  At N GPUs the paged cache is sharded by KV head (north_star):
every rank decodes Hkv/N KV heads of every request, no collective inside the timed region;
the NCCL all-gather of O is timed and reported separately (`allgather_us`).

`--impl reference`: the reference has no CPU implementation of this path (it is a CUDA library).
When oracle/_ref/liboracle_ref.so is present (the reference's own DecodePlan + decode + merge
templates compiled unmodified from its headers, see oracle/Makefile) and a GPU is visible, the
reference arm times THAT on the same workload: its plan + its kernels, nothing of ours on the
path.  Otherwise (or with --cpu-reference) it times the oracle port (oracle/attention_ref.py =
the reference test-suite's fp32 PyTorch attention) on the host cores over a bounded sample of
the workload, scaled to the full batch; that CPU figure is also what `cpu_baseline` reports.
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

PAGE = 16
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


def page_table(kv_lens, seed=0):
    pages = [-(-l // PAGE) for l in kv_lens]
    total = sum(pages)
    g = torch.Generator().manual_seed(seed)
    indices = torch.randperm(total, generator=g).to(torch.int32)  # shuffled: exercises the gather
    indptr = torch.zeros(len(kv_lens) + 1, dtype=torch.int32)
    indptr[1:] = torch.tensor(np.cumsum(pages), dtype=torch.int32)
    last = torch.tensor([((l - 1) % PAGE + 1) for l in kv_lens], dtype=torch.int32)
    return indptr, indices, last, total


# --------------------------------------------------------------------------- clocks sampler
class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop_evt = threading.Event()
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
            time.sleep(0.002)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=1)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


# --------------------------------------------------------------------------- CPU baseline / reference arm
def cpu_reference_time(wl, sample_requests: int, reps: int):
    """Time the oracle port on the host cores on `sample_requests` requests of the workload
    (full KV length, all heads) and scale to the whole batch.  Returns (us_per_full_step, info)."""
    from oracle import attention_ref as orc

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    kv_lens = wl["kv_lens"][:sample_requests]
    hq, hkv = wl["hq"], wl["hkv"]
    indptr, indices, last, total = page_table(kv_lens, seed=1)
    g = torch.Generator().manual_seed(0)
    dt = torch.bfloat16
    cache = torch.randn(total, 2, PAGE, hkv, HEAD_DIM, generator=g).to(dt)
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


def run_reference_gpu(args, wl, so_path):
    """The reference's own plan + CUDA kernels (oracle/_ref) on one GPU, whole workload."""
    import ctypes as C

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    ref = C.CDLL(str(so_path))
    ref.ref_last_error.restype = C.c_char_p
    kv_lens, hq, hkv = wl["kv_lens"], wl["hq"], wl["hkv"]
    bs = len(kv_lens)
    kv_dtype = torch.float8_e4m3fn if wl["kv_dtype"] == "e4m3" else torch.bfloat16
    kv_es = 1 if wl["kv_dtype"] == "e4m3" else 2
    indptr, indices, last, total_pages = page_table(kv_lens)
    cache_bytes = total_pages * 2 * PAGE * hkv * HEAD_DIM * kv_es
    n_copies = max(1, min(8, math.ceil((512 << 20) / cache_bytes)))
    torch.manual_seed(1234)
    caches = []
    for _ in range(n_copies):
        c = torch.randn(total_pages, 2, PAGE, hkv, HEAD_DIM, device=dev, dtype=torch.bfloat16)
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

    def call(mode, i=0):
        k, v = caches[i % n_copies][:, 0], caches[i % n_copies][:, 1]
        rc = ref.ref_batch_decode_mode(
            vp(fws.data_ptr()), C.c_size_t(fws.numel()), vp(iws.data_ptr()), vp(pws.data_ptr()),
            C.c_size_t(iws.numel()), vp(indptr.data_ptr()), bs, hq, hkv, HEAD_DIM, PAGE, 1, kvd, 0, 0,
            vp(q.data_ptr()), C.c_int64(q.stride(0)), C.c_int64(q.stride(1)), vp(k.data_ptr()),
            vp(v.data_ptr()), C.c_int64(k.stride(0)), C.c_int64(k.stride(1)), C.c_int64(k.stride(2)),
            vp(tbl_dev[1].data_ptr()), vp(tbl_dev[0].data_ptr()), vp(tbl_dev[2].data_ptr()),
            vp(out.data_ptr()), None, C.c_float(1.0 / math.sqrt(HEAD_DIM)), vp(stream),
            vp(info.data_ptr()), mode)
        if rc != 0:
            raise RuntimeError(f"reference shim failed: {ref.ref_last_error().decode()}")

    call(1)
    for i in range(max(args.warmup, 30)):
        call(2, i)
    torch.cuda.synchronize()
    sampler = ClockSampler(0)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        call(2, i)
    e1.record()
    torch.cuda.synchronize()
    clocks = sampler.stop()
    us_step = e0.elapsed_time(e1) * 1e3 / args.steps
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
    e2e_us = float(np.median(t_e2e[2:])) * 1e6
    h2d = q_host.numel() * 2 + sum(t.numel() for t in tbl_host) * 4
    cpu_us, cpu_info = cpu_reference_time(wl, sample_requests=1, reps=1)
    line = {
        "impl": "reference", "metric": "paged_decode_us_per_step", "value": us_step, "unit": "us",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": us_step / 1e3,
        "higher_is_better": False, "scaling": "strong", "vs_baseline": None,
        "dtype": "bf16" if kv_es == 2 else "bf16 q/o, e4m3 kv", "data": "synthetic",
        "tokens_per_s": bs / (us_step * 1e-6),
        "config": {"workload": f"{args.workload}: {wl['desc']}", "page_table": "shuffled",
                   "reference": "the reference's own DecodePlan + BatchDecodeWithPagedKVCache (+ merge) "
                                "templates, compiled unmodified from its headers into oracle/_ref "
                                "(the reference has no CPU implementation); one GPU, unsharded",
                   "launches_per_step": int(info[10])},
        "cpu_baseline": {"value": cpu_us, "unit": "us", **cpu_info},
        "e2e": {"value": e2e_us, "unit": "us", "h2d_bytes_per_step": h2d,
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
    us, info = cpu_reference_time(wl, sample_requests=2, reps=max(1, min(args.steps, 30)))
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


# --------------------------------------------------------------------------- GPU arm
def run_ours(args):
    import torch.distributed as dist

    import flashinfer_b200 as fi
    from flashinfer_b200 import _lib

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
    kv_lens, hq_all, hkv_all = wl["kv_lens"], wl["hq"], wl["hkv"]
    from flashinfer_b200 import sharding

    try:
        # --shard-of N (development aid): time rank 0's slice of an N-way sharding on this one GPU
        shard = sharding.head_shard(hq_all, hkv_all, rank, world) if args.shard_of <= 1 else \
            sharding.head_shard(hq_all, hkv_all, 0, args.shard_of)
    except ValueError as e:
        raise SystemExit(str(e))
    hq, hkv = shard.num_qo_heads, shard.num_kv_heads
    bs = len(kv_lens)
    kv_dtype = torch.float8_e4m3fn if wl["kv_dtype"] == "e4m3" else torch.bfloat16
    kv_es = 1 if wl["kv_dtype"] == "e4m3" else 2
    q_dtype = torch.bfloat16
    indptr, indices, last, total_pages = page_table(kv_lens)
    indptr, indices, last = indptr.pin_memory(), indices.pin_memory(), last.pin_memory()

    # ---- synthetic inputs, resident in HBM.  Distinct cache copies are rotated when one
    # copy would fit in the 126 MB L2, so every step streams from HBM.
    cache_bytes = total_pages * 2 * PAGE * hkv * HEAD_DIM * kv_es
    n_copies = max(1, min(8, math.ceil((512 << 20) / cache_bytes)))
    torch.manual_seed(1234 + rank)
    caches = []
    for _ in range(n_copies):
        shape = (total_pages, 2, PAGE, hkv, HEAD_DIM) if args.kv_layout == "NHD" else \
            (total_pages, 2, hkv, PAGE, HEAD_DIM)
        c = torch.randn(*shape, device=dev, dtype=torch.bfloat16)
        caches.append(c.to(kv_dtype) if kv_es == 1 else c)
        del c
    k_scale = v_scale = None
    if kv_es == 1:
        k_scale, v_scale = 0.5, 0.5  # per-tensor scales (values are synthetic)
    q_host = torch.randn(bs, hq, HEAD_DIM).to(q_dtype).pin_memory()
    o_host = torch.empty(bs, hq, HEAD_DIM, dtype=q_dtype).pin_memory()
    q = q_host.to(dev)
    out = torch.empty_like(q)
    ws = torch.zeros(128 << 20, dtype=torch.uint8, device=dev)
    wrapper = fi.BatchDecodeWithPagedKVCacheWrapper(ws, args.kv_layout)
    lib = _lib.load()

    def plan():
        wrapper.plan(indptr, indices, last, hq, hkv, HEAD_DIM, PAGE, q_data_type=q_dtype,
                     kv_data_type=kv_dtype)

    def step(i):
        wrapper.run(q, caches[i % n_copies], k_scale=k_scale, v_scale=v_scale, out=out)

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    plan()
    # W untimed warm-up steps as asked, topped up to 30: the first launches after an idle GPU run at
    # ramping clocks (a 20-step timed region right after 3 warm-up steps reads 168 us instead of 162.6)
    for i in range(max(args.warmup, 30)):
        step(i)
    sync_all()

    # ---- device-resident timing: K steps between two events on the launching stream
    n0 = lib.fi_launch_count()
    sampler = ClockSampler(local_rank)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sync_all()
    e0.record()
    for i in range(args.steps):
        step(i)
    e1.record()
    sync_all()
    clocks = sampler.stop()
    launches = lib.fi_launch_count() - n0
    ms_total = e0.elapsed_time(e1)

    # ---- the same kernel timed one launch at a time (L2 flushed by reading a 256 MB buffer the host
    # enqueues ahead, so the launch is already queued when the first event fires): reported as
    # roofline.kernel_us_isolated; the roofline itself uses the average over the timed region
    flush = torch.zeros(64 << 20, dtype=torch.float32, device=dev)  # 256 MB, twice the L2
    per = []
    for i in range(min(args.steps, 50)):
        flush.sum()  # READ 256 MB: the L2 now holds someone else's clean lines, as between a model's GEMMs
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        step(i)
        b.record()
        b.synchronize()
        per.append(a.elapsed_time(b))
    kern_ms = float(np.median(per))
    del flush

    # ---- end to end through the public API with HOST buffers: plan (page table from host
    # memory) + q H2D from pinned memory + run + o D2H, every step
    sync_all()
    t_e2e = []
    for i in range(max(3, min(args.steps, 20)) + 2):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        q.copy_(q_host, non_blocking=True)  # enqueued first: the copy runs while the host plans
        plan()
        step(i)
        o_host.copy_(out, non_blocking=True)
        torch.cuda.synchronize()
        t_e2e.append(time.perf_counter() - t0)
    e2e_us = float(np.median(t_e2e[2:])) * 1e6
    h2d = q_host.numel() * 2 + (indptr.numel() + indices.numel() + last.numel()) * 4
    d2h = o_host.numel() * 2

    # ---- the only collective: all-gather of O across head shards, timed separately
    ag_us = None
    if world > 1:
        for _ in range(5):
            full = sharding.gather_heads(out, world)
        assert full.shape == (bs, hq_all, HEAD_DIM)
        sync_all()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(20):
            sharding.gather_heads(out, world)
        b.record()
        sync_all()
        ag_us = a.elapsed_time(b) / 20 * 1e3

    # ---- max over ranks
    vals = torch.tensor([ms_total, kern_ms, e2e_us, ag_us or 0.0], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(vals, op=dist.ReduceOp.MAX)
    ms_total, kern_ms, e2e_us, ag_max = [float(x) for x in vals.tolist()]
    if rank == 0:
        emit(args, wl, world, ms_total, kern_ms, e2e_us, ag_max if world > 1 else None, h2d, d2h,
             launches, clocks, hq, hkv, kv_es, n_copies, cache_bytes)
    if world > 1:
        dist.destroy_process_group()


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


def emit(args, wl, world, ms_total, kern_ms, e2e_us, ag_us, h2d, d2h, launches, clocks, hq, hkv,
         kv_es, n_copies, cache_bytes):
    bs = len(wl["kv_lens"])
    us_step = ms_total * 1e3 / args.steps
    bytes_rank = algorithmic_bytes(wl["kv_lens"], hq, hkv, kv_es)  # one launch on one rank
    peak, peak_src = hbm_peak_gbs()
    # one launch per step and nothing else in the timed region: the kernel's average launch
    # duration is the step time (events on the launching stream around the K steps)
    achieved = bytes_rank / (us_step * 1e-6) / 1e9
    line = {
        "metric": "paged_decode_us_per_step", "value": us_step, "unit": "us", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "warmup_run": max(args.warmup, 30),
        "ms_per_step": us_step / 1e3,
        "higher_is_better": False, "scaling": "strong", "vs_baseline": None,
        "dtype": "bf16" if kv_es == 2 else "bf16 q/o, e4m3 kv", "data": "synthetic",
        "tokens_per_s": bs / (us_step * 1e-6),
        "hbm_tbps_whole_job": bytes_rank * world / (us_step * 1e-6) / 1e12,
        "config": {"workload": f"{args.workload}: {wl['desc']}" +
                               (" [cache laid out HND for this run]" if args.kv_layout != "NHD" else ""),
                   "sharding": f"kv-head x{world}" if args.shard_of <= 1 else
                               f"ONE rank of a kv-head x{args.shard_of} sharding (not the whole job)",
                   "heads_per_rank": [hq, hkv], "page_table": "shuffled",
                   "l2": f"KV per rank {cache_bytes / 2**20:.0f} MiB > L2; {n_copies} cache "
                         f"cop{'y' if n_copies == 1 else 'ies'} rotated"},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "frac_of_8tbps": achieved / 8000.0,
                     "traffic": ncu_traffic_per_launch(args.workload, world), "peak_source": peak_src,
                     "kernel": "paged_decode_kernel / paged_decode_tc5_kernel (split-KV merge fused), "
                               "one launch per run()",
                     "kernel_us": us_step, "kernel_us_isolated": kern_ms * 1e3,
                     "algorithmic_bytes_per_launch": bytes_rank},
        "e2e": {"value": e2e_us, "unit": "us", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "what": "q H2D (pinned, enqueued first so it overlaps the host planner) + plan (page table "
                        "in pinned host memory -> work list + indices H2D) + run + o D2H, per step"},
        "gpu_launches": int(launches), "clocks": clocks,
    }
    if ag_us is not None:
        line["allgather_us"] = ag_us
    if world == 1 and not args.no_cpu:
        # a quarter of the C2 batch (~0.2 s per pass on 16 cores); 2 requests of the heavier workloads
        us, info = cpu_reference_time(wl, sample_requests=16 if args.workload == "c2" else 2, reps=3)
        line["cpu_baseline"] = {"value": us, "unit": "us", **info}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=["c2", "c3", "c4", "c5"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
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
