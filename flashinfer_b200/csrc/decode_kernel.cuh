// decode_kernel.cuh — persistent paged-KV decode attention kernel for sm_100a.
//
// Replaces BatchDecodeWithPagedKVCacheKernel / SingleDecodeWithKVCacheKernel
// (reference include/flashinfer/attention/decode.cuh:217-368,393-617).  Design:
//   * one persistent CTA per (SM x occupancy); each CTA walks a planner-made list of
//     segments = contiguous KV-tile ranges of one (request, kv_head) unit;
//   * producer warp(s) stream 64-token K/V tiles through a multi-stage shared-memory
//     ring guarded by full/empty mbarriers: per-page TMA boxes (128B-swizzled) when
//     the page geometry allows, else 16-byte cp.async gathers through the page table;
//   * 4 consumer warps, 16 tokens each per tile.  All G (<= 8) query heads of the GQA
//     group ride in the N=8 dimension of mma.sync.m16n8k16:
//        S^T[token, head] = K[token, :] . Q[head, :]      (A = K tile via ldmatrix)
//        O^T[d, head]    += V^T[d, token] . P^T[token, head] (A = V via ldmatrix.trans,
//                                                          B = P via movmatrix)
//     online softmax in base 2 with warp-shuffle max reductions over the token lanes;
//   * FP8 (e4m3/e5m2) KV is dequantised in registers to f16 right after ldmatrix;
//   * at a segment end the 4 warp states are merged through shared memory and either
//     the final o/lse (whole unit) or a normalised fp32 partial (split unit) is written.
// LSE is base 2 and the "-inf" sentinel is the finite -5e4 (reference state.cuh:45,
// math.cuh:32).
#pragma once
#include <cuda.h>

#include <type_traits>

#include "decode_params.h"
#include "ptx.cuh"

namespace fib200 {

// Phase timestamps for latency work (scripts/trace_probe.py): compiled only into the
// -DFI_B200_TRACE build (libfi_b200_trace.so), never into the shipped library.
#ifdef FI_B200_TRACE
constexpr int kTraceCtas = 1024, kTraceSlots = 16;
__device__ unsigned long long g_fi_trace[kTraceCtas * kTraceSlots];
__device__ __forceinline__ void fi_trace(int slot) {
  if (blockIdx.x < kTraceCtas) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    g_fi_trace[blockIdx.x * kTraceSlots + slot] = t;
  }
}
__device__ __forceinline__ void fi_trace_smid() {
  if (blockIdx.x < kTraceCtas) {
    unsigned id;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(id));
    g_fi_trace[blockIdx.x * kTraceSlots + 15] = id + 1;
  }
}
#define FI_TRACE(slot) fi_trace(slot)
#define FI_TRACE_IF(cond, slot) do { if (cond) fi_trace(slot); } while (0)
#else
#define FI_TRACE(slot)
#define FI_TRACE_IF(cond, slot)
#endif

// cp.async gather loader: FOUR producer warps = one warpgroup, one warp on each of the SM's four
// sub-partitions (cp.async goes through each sub-partition's load/store instruction queue; with two
// producer warps half of the queues stayed idle).  Two 256-thread CTAs per SM fit the register file
// because the producer warpgroup hands registers to the consumer warpgroup (setmaxnreg).
constexpr int kCpAsyncProducerThreads = 128;
constexpr int kCpAsyncConsumerRegs = 176, kCpAsyncProducerRegs = 80;  // 128 * (176 + 80) = 32768

template <int KVF, int HD>
struct TileGeom {
  static constexpr int ES = (KVF == 0) ? 2 : 1;          // bytes per KV element
  static constexpr int TT = tile_tokens(KVF);            // tokens per stage
  static constexpr int TPW = TT / kConsumerWarps;        // tokens per consumer warp per stage
  static constexpr int ROW_BYTES = HD * ES;              // one token row of K (or V)
  static constexpr int NP = ROW_BYTES / 128;             // 128-byte swizzle panels per row
  static constexpr int CHUNKS = ROW_BYTES / 16;          // 16-byte chunks per row
  static constexpr int PANEL_BYTES = TT * 128;           // 8 / 16 KiB
  static constexpr int KV_BYTES = NP * PANEL_BYTES;      // K part of one stage
  static constexpr int STAGE_BYTES = 2 * KV_BYTES;       // K + V
  static constexpr int KSTEPS = HD / 16;                 // mma k-steps for QK == m-tiles for PV
  static_assert(ROW_BYTES % 128 == 0, "row must be a whole number of 128B panels");
};

// shared-memory carve-up (offsets from the 1024-aligned base)
template <int KVF, int HD>
struct SmemLayout {
  using G = TileGeom<KVF, HD>;
  static constexpr int SCRATCH_O = (kConsumerWarps - 1) * G::KSTEPS * 4 * 32 * 4;
  static constexpr int SCRATCH_ML = kScratchMlBytes;  // warp (m, l) pairs + merge flag; tcgen05 path: more
  static constexpr int SEG_BYTES = kSegCache * (int)sizeof(Segment);
  static constexpr int BAR_BYTES = 2 * kMaxStages * 8 + 128;  // ring + tcgen05-path barriers
  __host__ __device__ static constexpr int ring(int stages) { return stages * G::STAGE_BYTES; }
  __host__ __device__ static constexpr int total(int stages) {
    return 1024 + ring(stages) + SCRATCH_O + SCRATCH_ML + SEG_BYTES + BAR_BYTES;
  }
};

// byte offset of (row, 16B-chunk) inside the K (or V) part of a stage, 128B swizzle — the layout of
// the tcgen05 kernel (its MMA descriptors want the 128 K rows of a tile contiguous): K part, V part
template <int TT>
__device__ __forceinline__ uint32_t swz(int row, int chunk) {
  return (uint32_t)((chunk >> 3) * (TT * 128) + row * 128 +
                    (((chunk & 7) ^ (row & 7)) << 4));
}

// Layout of the mma.sync kernel: K and V of the same token block sit next to each other, so that ONE
// 5-D TMA box {128 B, rows_per_box tokens, 1 head, K|V, 1 page} fills both — half as many TMA
// instructions per tile as a K box and a V box (the TMA unit's per-box cost, not DRAM, bounds small
// boxes: C2 at page 8 ran at 227 us vs 162 us at page 16 and 158 us at page 32 with K / V boxes):
//   stage = [128-byte panel pn][token block j of rpb rows][K rows | V rows][row][128 B, swizzled]
// rmask = ~(rpb - 1); the V row is vdelta = rpb * 128 bytes after its K row.  With rpb = TT (the
// cp.async loader) this degenerates to [panel][K | V][row].
template <int TT>
struct StageRow {
  uint32_t base, r7;
  __device__ __forceinline__ StageRow(int row, int rmask)
      : base((uint32_t)(row + (row & rmask)) * 128u), r7((uint32_t)row & 7u) {}
  __device__ __forceinline__ uint32_t at(int chunk) const {  // K; V: + vdelta
    return (uint32_t)(chunk >> 3) * (2u * TT * 128u) + base + ((((uint32_t)chunk & 7u) ^ r7) << 4);
  }
};

// ---- segment queue ---------------------------------------------------------------
// A CTA works through a run-time sequence of segments: those of item blockIdx.x, then those of
// every item it draws from the plan's ticket counter.  One warp (the "feeder": the TMA producer
// warp, or warp 0 of the cp.async producers) draws tickets, fetches item heads / list segments
// and publishes them into a shared-memory ring of kSegCache Segment records; all other roles
// read the ring.  Control words (ints in shared memory):
constexpr int kQPub = 0;    // segments published so far, | kQFinal once nothing more will come
constexpr int kQDone0 = 1;  // consumer group: ring entries below this index are no longer read
constexpr int kQDone1 = 2;  // second reader (cp.async producer warp 1 / tcgen05 issuer)
constexpr int kQPeek = 3;   // tcgen05 softmax: "segment k + 1 is already published" broadcast
constexpr int kQDone2 = 4;  // cp.async producer warps 2 and 3
constexpr int kQDone3 = 5;
constexpr int kQFinal = 1 << 30;
constexpr int kQNoReader = 0x3fffffff;
constexpr int kQueueCtlFloat = 160;  // offset of the control words in the warp scratch (floats)
constexpr int kMergeDescFloat = 168; // {go, request, kv_head, slot_begin, count} of a posted merge
constexpr int kHeadPagesFloat = 176; // kItemInlinePages page ids of item blockIdx.x's first tiles
static_assert((kHeadPagesFloat + kItemInlinePages) * 4 <= kScratchMlBytes, "inline pages do not fit the scratch");

// In the implicit mode (single request, header == nullptr) the sequence is static and in closed
// form: nothing is drawn, the ring holds the first kSegCache segments and later ones are recomputed.
struct SegSource {
  const DecodeParams* p;
  Segment* cache;
  volatile int* ctl;
  bool implicit;
  int first, head_count, num_items;  // explicit: head of item blockIdx.x
  int num_pages;                     // explicit: page ids of the item's first tiles held in `pages`
  const int* pages;                  //   (shared memory), 0 when plan() did not have the table
  int count;                         // implicit: segments of this CTA
  long long imp_begin, imp_end;      // implicit: stream range of this CTA
  int tt;                            // tokens per tile
};

// closed-form segment k of this CTA (implicit mode only)
__device__ __forceinline__ Segment seg_make(const SegSource& s, int k) {
  const DecodeParams& p = *s.p;
  Segment g;
  g.request = 0; g.merge_unit = -1; g.page_base = 0;
  g.merge_count = 0; g.merge_slot = 0;
  g.tok_lo = p.imp_tok_lo; g.tok_hi = p.imp_tok_hi; g.kv_len = p.imp_tok_hi;
  const int nt = p.imp_tiles;
  if (nt == 0) {
    g.kv_head = k; g.tile_begin = g.tile_end = 0; g.slot = -1;
  } else {
    const int tb0 = p.imp_tok_lo / s.tt;  // first tile of the window
    const long long u = s.imp_begin / nt + k;
    const long long ub = u * nt, ue = ub + nt;
    const long long b = s.imp_begin > ub ? s.imp_begin : ub;
    const long long e = s.imp_end < ue ? s.imp_end : ue;
    g.kv_head = (int)u;
    g.tile_begin = tb0 + (int)(b - ub);
    g.tile_end = tb0 + (int)(e - ub);
    const bool whole = (b == ub) && (e == ue);
    g.slot = whole ? -1 : (int)(blockIdx.x + u);
  }
  return g;
}

// implicit mode: this CTA's share of the closed-form uniform split
__device__ __forceinline__ void seg_source_implicit_range(SegSource& s) {
  const DecodeParams& p = *s.p;
  const long long T = (long long)p.imp_units * p.imp_tiles;
  const long long C = gridDim.x;
  s.count = 0;
  s.imp_begin = s.imp_end = 0;
  if (T == 0) {
    s.count = (blockIdx.x == 0) ? p.imp_units : 0;  // empty KV: zero-tile segments
  } else {
    s.imp_begin = ((long long)blockIdx.x * T) / C;
    s.imp_end = ((long long)(blockIdx.x + 1) * T) / C;
    if (s.imp_end > s.imp_begin)
      s.count = (int)((s.imp_end - 1) / p.imp_tiles - s.imp_begin / p.imp_tiles) + 1;
  }
}

// Called by every thread of the CTA before the first __syncthreads(): reads the head of item
// blockIdx.x (one global round trip), publishes its inline segments and sets up the control words.
// has_reader1: the kernel has a second ring reader that reports progress in kQDone1.
__device__ __forceinline__ SegSource seg_source_init(const DecodeParams& p, int tt, Segment* cache,
                                                     float* scratch, bool has_reader1,
                                                     bool has_reader23 = false) {
  SegSource s;
  s.p = &p;
  s.cache = cache;
  s.ctl = reinterpret_cast<volatile int*>(scratch + kQueueCtlFloat);
  s.tt = tt;
  s.implicit = p.header == nullptr;
  s.first = s.head_count = s.num_items = s.count = 0;
  s.imp_begin = s.imp_end = 0;
  s.num_pages = 0;
  s.pages = reinterpret_cast<const int*>(scratch + kHeadPagesFloat);
  const int tid = threadIdx.x;
  int pub0;
  if (!s.implicit) {
    const ItemHead* head = p.item_heads + blockIdx.x;  // heads exist for every launched CTA
    const uint64_t keep = l2_policy_evict_last();
    const int4 h = (p.l2_hints & 2) ? ldg_int4_hint(reinterpret_cast<const int4*>(head), keep)
                              : __ldg(reinterpret_cast<const int4*>(head));
    s.first = h.x;
    s.head_count = h.y;
    s.num_items = h.z;
    s.num_pages = h.w;
    pub0 = min(s.head_count, kItemInlineSegs);
    if (tid < 3 * pub0)
      reinterpret_cast<int4*>(cache)[tid] =
          (p.l2_hints & 2) ? ldg_int4_hint(reinterpret_cast<const int4*>(head->seg) + tid, keep)
                     : __ldg(reinterpret_cast<const int4*>(head->seg) + tid);
    else if (tid >= 32 && tid < 32 + kItemInlinePages / 4 && h.w > 0)
      reinterpret_cast<int4*>(scratch + kHeadPagesFloat)[tid - 32] =
          (p.l2_hints & 2) ? ldg_int4_hint(reinterpret_cast<const int4*>(head->pages) + (tid - 32), keep)
                     : __ldg(reinterpret_cast<const int4*>(head->pages) + (tid - 32));
  } else {
    seg_source_implicit_range(s);
    for (int i = tid; i < min(s.count, kSegCache); i += blockDim.x) cache[i] = seg_make(s, i);
    pub0 = s.count | kQFinal;
  }
  if (tid == 0) {
    s.ctl[kQPub] = pub0;
    s.ctl[kQDone0] = 0;
    s.ctl[kQDone1] = has_reader1 ? 0 : kQNoReader;
    s.ctl[kQDone2] = s.ctl[kQDone3] = has_reader23 ? 0 : kQNoReader;
    s.ctl[kQPeek] = 0;
  }
  return s;
}

// segment k, which the caller knows to be published
__device__ __forceinline__ Segment seg_read(const SegSource& s, int k) {
  if (s.implicit && k >= kSegCache) return seg_make(s, k);
  return s.cache[k & (kSegCache - 1)];
}

// is segment k published yet?  (non-blocking)
__device__ __forceinline__ bool seg_peek(const SegSource& s, int k) {
  if (s.implicit) return k < s.count;
  const bool yes = k < (s.ctl[kQPub] & (kQFinal - 1));
  __threadfence_block();
  return yes;
}

// wait for segment k; false when the CTA's sequence ended before it
__device__ __forceinline__ bool seg_fetch(const SegSource& s, int k, Segment& out) {
  if (s.implicit) {
    if (k >= s.count) return false;
  } else {
    for (;;) {
      const int v = s.ctl[kQPub];
      if (k < (v & (kQFinal - 1))) break;
      if (v & kQFinal) return false;
      __nanosleep(40);
    }
    __threadfence_block();
  }
  out = seg_read(s, k);
  return true;
}

// ---- feeder: draws items and publishes their segments (one warp, converged) -------------
// One latency-bound step per call — draw a ticket / load the item head or 8 list segments /
// publish — so that called once per tile from the producer loop the global round trips hide
// behind the tile traffic.  A new item is drawn when at most sched_lookahead published tiles
// are still to be issued: late enough that a slow CTA does not hoard tail items.
struct Feeder {
  int pub, fin, item_first, item_count, item_next, num_items, tiles_pub, phase, ticket, nload;
  int4 v;
};

// is there nothing at all for this CTA to do?  (static plans: an empty head; plans with more items
// than CTAs: every CTA draws tickets, an empty first item does not excuse it)
__device__ __forceinline__ bool seg_source_idle(const SegSource& s) {
  if (s.implicit) return s.count == 0;
  return s.head_count == 0 && s.num_items <= (int)gridDim.x;
}

__device__ __forceinline__ void feeder_init(Feeder& f, const SegSource& s) {
  f.fin = s.implicit ? 1 : 0;
  f.pub = s.implicit ? s.count : min(s.head_count, kItemInlineSegs);
  f.item_first = s.first;
  f.item_count = s.implicit ? 0 : s.head_count;
  f.item_next = s.implicit ? 0 : f.pub;
  f.num_items = s.num_items;
  f.tiles_pub = 0;
  if (!s.implicit)
    for (int i = 0; i < f.pub; ++i) f.tiles_pub += s.cache[i].tile_end - s.cache[i].tile_begin;
  f.phase = 0;
  f.ticket = 0;
  f.nload = 0;
  f.v = make_int4(0, 0, 0, 0);
}

__device__ __forceinline__ void feeder_step(Feeder& f, const SegSource& s, int cur_k,
                                            int tiles_issued, int lane) {
  if (f.fin) return;
  const DecodeParams& p = *s.p;
  if (f.phase >= 2) {  // loaded records -> ring, publish
    int n, base;
    if (f.phase == 2) {  // lane 0: head, lanes 1..9: inline segments
      f.item_first = __shfl_sync(0xffffffffu, f.v.x, 0);
      f.item_count = __shfl_sync(0xffffffffu, f.v.y, 0);
      n = min(f.item_count, kItemInlineSegs);
      base = 1;
      f.item_next = n;
    } else {  // lanes 0..3n-1: list segments
      n = f.nload;
      base = 0;
      f.item_next += n;
    }
    const int li = lane - base;
    const bool mine = li >= 0 && li < 3 * n;
    const int j = li / 3, part = li - 3 * j;
    if (mine) reinterpret_cast<int4*>(s.cache + ((f.pub + j) & (kSegCache - 1)))[part] = f.v;
    f.tiles_pub += __reduce_add_sync(0xffffffffu, (mine && part == 0) ? f.v.w - f.v.z : 0);
    f.pub += n;
    __syncwarp();
    __threadfence_block();
    if (lane == 0) s.ctl[kQPub] = f.pub;
    f.phase = 0;
    return;
  }
  if (f.phase == 1) {  // ticket -> item head
    const int t = __shfl_sync(0xffffffffu, f.ticket, 0);
    const int item = t + (int)gridDim.x;
    // exactly num_items tickets are drawn per launch (every item past the static ones once,
    // plus one failing draw per active CTA): the last one re-arms the counter
    if (t == f.num_items - 1 && lane == 0) *p.work_counter = 0;
    if (item >= f.num_items) {
      f.fin = 1;
      if (lane == 0) s.ctl[kQPub] = f.pub | kQFinal;
      f.phase = 0;
      return;
    }
    if (lane < 1 + 3 * kItemInlineSegs)
      f.v = __ldg(reinterpret_cast<const int4*>(p.item_heads + item) + lane);
    f.phase = 2;
    return;
  }
  // idle: is there room in the ring, and is it time to fetch more?
  const int done = min(min(s.ctl[kQDone0], s.ctl[kQDone1]), min(s.ctl[kQDone2], s.ctl[kQDone3]));
  const bool space = f.pub + kSegRefill <= done + kSegCache;
  if (f.item_next < f.item_count) {  // rest of the current item, from the list
    if (space && f.pub - cur_k < kSegRefill) {
      const int n = min(kSegRefill, f.item_count - f.item_next);
      const Segment* list =
          reinterpret_cast<const Segment*>(p.item_heads + max(f.num_items, (int)gridDim.x));
      if (lane < 3 * n)
        f.v = __ldg(reinterpret_cast<const int4*>(list + f.item_first + f.item_next) + lane);
      f.nload = n;
      f.phase = 3;
    }
  } else if (f.num_items <= (int)gridDim.x) {  // static plan: item blockIdx.x was all there is
    f.fin = 1;
    if (lane == 0) s.ctl[kQPub] = f.pub | kQFinal;
  } else if (space && f.tiles_pub - tiles_issued <= p.sched_lookahead) {
    if (lane == 0) f.ticket = atomicAdd(p.work_counter, 1);
    f.phase = 1;
  }
}

// feeder-side: make sure segment k is published (false: the sequence ended before it)
__device__ __forceinline__ bool feeder_await(Feeder& f, const SegSource& s, int k, int tiles_issued,
                                             int lane) {
  while (k >= f.pub && !f.fin) feeder_step(f, s, k, tiles_issued, lane);
  return k < f.pub;
}

// ---- producer: per-page TMA boxes -------------------------------------------
// Runs in one warp (all 32 lanes converged).  Box i of a tile is issued by lane i%32:
//   i = (kv * NP + panel) * BPT + j,  j = token-block of rows_per_box rows.
// (A variant that also pulled tiles beyond the ring into L2 with cp.async.bulk.prefetch.tensor
// was measured 1.5x SLOWER on B200 — the prefetch boxes compete with the demand boxes — and
// was removed; see DESIGN.md.)
// PAIRED (the mma.sync kernel): the StageRow layout; with p.tma_paired the K and V panels of a
// (page, token block) come in ONE 5-D box through `kmap`, else as a K box and a V box.
// HEADPAIR (the tcgen05 kernel on a head-pair plan, fp8 KV): a tile is G::TT / 2 tokens of TWO
// adjacent KV heads; kmap / vmap are 5-D (128 B, 8 tokens, head, token / 8, page) and one box
// {128, 8, 2, rows_per_box / 8, 1} brings rows_per_box tokens of both heads as 8-row groups in the
// order (token / 8, head) — every 1 KiB group is one swizzle atom of the MMA descriptors, and a
// token's two 128-byte rows (adjacent in an NHD cache) are requested together.
template <int KVF, int HD, bool PAIRED = false, bool HEADPAIR = false>
__device__ __forceinline__ void produce_tma(const DecodeParams& p, const SegSource& src,
                                            const CUtensorMap* kmap, const CUtensorMap* vmap,
                                            uint32_t ring, uint32_t full_bar, uint32_t empty_bar) {
  using G = TileGeom<KVF, HD>;
  static_assert(!HEADPAIR || (!PAIRED && G::NP == 1), "head pairs: 128-byte rows, K and V boxes");
  constexpr int TT = HEADPAIR ? G::TT / 2 : G::TT;
  const int lane = threadIdx.x & 31;
  const int rpb = p.rows_per_box;
  const int bpt = TT / rpb;  // boxes per tile along tokens: a power of two <= 16
  const bool one_box = PAIRED && p.tma_paired != 0;
  const int nbox = (one_box ? 1 : 2) * G::NP * bpt;
  const int j = lane % bpt;  // same j for every i = lane + 32 k since bpt | 32
  const uint32_t box_bytes = (uint32_t)rpb * (HEADPAIR ? 256u : 128u);

  const bool hints = (p.l2_hints & 1) != 0, keep_meta = (p.l2_hints & 2) != 0;
  const uint64_t pol_stream = l2_policy_evict_first(), pol_keep = l2_policy_evict_last();

  // Programmatic dependent launch: the work list and the page table were fixed at plan() time (host
  // memcpys ordered before every later kernel), so the first tile's page indices are requested
  // BEFORE the dependency wait and are in flight while the upstream kernel drains; the KV cache
  // itself (append_paged_kv_cache may be that upstream kernel) is only touched after the wait.
  // Plans with drawn items touch the ticket counter, which the previous launch re-arms: wait first.
  // A page table the caller keeps on the device (fast_decode_plan, device tensors given to plan())
  // may be rewritten by the upstream kernel: then nothing is read from it before the wait.
  bool waited = p.enable_pdl == 0;
  if (!waited && (src.num_items > (int)gridDim.x || !p.indices_early)) { pdl_wait(); waited = true; }

  Feeder f;
  feeder_init(f, src);
  Segment sg;
  int k = 0, t = 0, t_end = 0, tiles_issued = 0, mem_head = 0;
  int page0 = 0, npage0 = 0;  // segment 0 only: its first pages are in the item head (shared memory)
  int stage = 0;
  uint32_t phase = 0;
  [[maybe_unused]] bool first_tile = true;
  for (;;) {
    feeder_step(f, src, k, tiles_issued, lane);
    if (t >= t_end) {  // next segment that has tiles
      bool more;
      for (;;) {
        more = feeder_await(f, src, k, tiles_issued, lane);
        if (!more) break;
        sg = seg_read(src, k);
        if (sg.tile_end > sg.tile_begin) break;
        ++k;
      }
      if (!more) break;
      t = sg.tile_begin;
      t_end = sg.tile_end;
      mem_head = HEADPAIR ? 2 * sg.kv_head : sg.kv_head / p.kv_head_div;
      npage0 = k == 0 ? src.num_pages : 0;
      page0 = (sg.tile_begin * TT) >> p.page_shift;
    }
    // boxes of tile t: page index + first row for this lane's token block j
    const int tok0 = t * TT;
    const int rem = sg.tok_hi - tok0;  // > 0 by construction
    const int nvb = rem >= TT ? bpt : (rem + rpb - 1) / rpb;
    const int tok = tok0 + j * rpb;
    int page = 0;
    if (j < nvb) {
      const int pit = tok >> p.page_shift;
      if (p.kv_indices == nullptr) page = sg.page_base + pit;
      else if ((unsigned)(pit - page0) < (unsigned)npage0) page = src.pages[pit - page0];
      else page = keep_meta ? ldg_s32_hint(p.kv_indices + sg.page_base + pit, pol_keep)
                        : __ldg(p.kv_indices + sg.page_base + pit);
    }
    const int row0 = tok & p.page_mask;
    if (!waited) { pdl_wait(); waited = true; }
    FI_TRACE_IF(lane == 0 && first_tile && page >= 0, 3);
    mbar_wait(empty_bar + 8 * stage, phase ^ 1);
    if (lane == 0) mbar_arrive_expect_tx(full_bar + 8 * stage, 2u * G::NP * nvb * box_bytes);
    __syncwarp();
    if (j < nvb) {
      const uint32_t sbase = ring + stage * G::STAGE_BYTES + (PAIRED ? 2 : 1) * j * box_bytes;
      const int c1 = p.tma_h_first ? mem_head : row0, c2 = p.tma_h_first ? row0 : mem_head;
      for (int i = lane; i < nbox; i += 32) {
        const int pn = (i / bpt) % G::NP;
        const int kv = i / (bpt * G::NP);
        if constexpr (HEADPAIR) {
          const uint32_t dst = sbase + kv * G::KV_BYTES;
          if (hints) tma_load_5d_hint(dst, kv ? vmap : kmap, full_bar + 8 * stage, 0, 0, mem_head, row0 >> 3, page, pol_stream);
          else tma_load_5d(dst, kv ? vmap : kmap, full_bar + 8 * stage, 0, 0, mem_head, row0 >> 3, page);
        } else if constexpr (PAIRED) {
          const uint32_t dst = sbase + pn * (2 * G::PANEL_BYTES) + kv * box_bytes;
          if (one_box) {
            if (hints) tma_load_5d_hint(dst, kmap, full_bar + 8 * stage, pn * (128 / G::ES), c1, c2, 0, page, pol_stream);
            else tma_load_5d(dst, kmap, full_bar + 8 * stage, pn * (128 / G::ES), c1, c2, 0, page);
          } else if (hints) {
            tma_load_4d_hint(dst, kv ? vmap : kmap, full_bar + 8 * stage, pn * (128 / G::ES), c1, c2, page, pol_stream);
          } else {
            tma_load_4d(dst, kv ? vmap : kmap, full_bar + 8 * stage, pn * (128 / G::ES), c1, c2, page);
          }
        } else if (hints) {
          tma_load_4d_hint(sbase + kv * G::KV_BYTES + pn * G::PANEL_BYTES, kv ? vmap : kmap,
                           full_bar + 8 * stage, pn * (128 / G::ES), c1, c2, page, pol_stream);
        } else {
          tma_load_4d(sbase + kv * G::KV_BYTES + pn * G::PANEL_BYTES, kv ? vmap : kmap,
                      full_bar + 8 * stage, pn * (128 / G::ES), c1, c2, page);
        }
      }
    }
    __syncwarp();
    FI_TRACE_IF(lane == 0 && first_tile, 4);
    first_tile = false;
    if (++stage == p.num_stages) { stage = 0; phase ^= 1; }
    ++tiles_issued;
    if (++t >= t_end) ++k;
  }
}

// ---- producer: 16-byte cp.async gathers (any page size / stride) -------------
// Runs in kCpAsyncProducerThreads threads (ptid = 0 ..): warp 0 is also the feeder, warps 1 .. 3 read
// the ring like consumers.  Rows past tok_hi are zero-filled.
// A thread owns one 16-byte chunk column of PASSES rows of every tile.  Two things bounded the first
// version of this loader at 2.1x .. 4.3x the TMA loader's time (C2: 342 us at page 16, 690 us at page
// 1, against 162 us): a dependent page-table load in front of every row's address, and ~40 integer
// instructions per row in two warps.  Now
//   * the page ids of a tile are fetched ONE TILE AHEAD, one table entry per lane (a tile touches at
//     most 64 pages), and handed to the rows by warp shuffles;
//   * everything about a row that does not depend on the tile (destination, offset inside the page)
//     is folded at compile time or hoisted, so a row costs one 64-bit multiply-add and two cp.async.
template <int KVF, int HD>
struct CpAsyncGeom {
  using G = TileGeom<KVF, HD>;
  static constexpr int CPR = G::CHUNKS;
  static constexpr int RPP = kCpAsyncProducerThreads / CPR;  // rows per pass
  static constexpr int PASSES = G::TT / RPP;
  static constexpr int SLOTS = (G::TT + 31) / 32;            // page-id registers per lane
  static_assert(kCpAsyncProducerThreads % CPR == 0 && G::TT % RPP == 0, "bad cp.async geometry");
};

// page ids of tile t of segment sg: slot i <-> the i-th page the tile touches, lane (i & 31) holds it
template <int KVF, int HD>
__device__ __forceinline__ void cpasync_tile_pages(int (&pg)[CpAsyncGeom<KVF, HD>::SLOTS],
                                                   const DecodeParams& p, const Segment& sg, int t,
                                                   int lane) {
  using G = TileGeom<KVF, HD>;
  const int tok0 = t * G::TT;
  const int pit0 = p.page_shift >= 0 ? tok0 >> p.page_shift : tok0 / p.page_size;
  const int last_tok = min(tok0 + G::TT, sg.tok_hi) - 1;  // >= tok0: a listed tile has tokens
  const int pit1 = p.page_shift >= 0 ? last_tok >> p.page_shift : last_tok / p.page_size;
#pragma unroll
  for (int i = 0; i < CpAsyncGeom<KVF, HD>::SLOTS; ++i) {
    const int pit = pit0 + 32 * i + lane;
    int page = 0;
    if (pit <= pit1) page = p.kv_indices ? __ldg(p.kv_indices + sg.page_base + pit) : sg.page_base + pit;
    pg[i] = page;
  }
}

template <int KVF, int HD>
__device__ __forceinline__ void produce_cpasync(const DecodeParams& p, const SegSource& src, int ptid,
                                                uint32_t ring, uint32_t full_bar,
                                                uint32_t empty_bar) {
  using G = TileGeom<KVF, HD>;
  using R = CpAsyncGeom<KVF, HD>;
  const int c = ptid % R::CPR;
  const int r0 = ptid / R::CPR;
  const int lane = ptid & 31;
  const bool feeds = ptid < 32;
  const char* kb = static_cast<const char*>(p.k);
  const char* vb = static_cast<const char*>(p.v);
  // page sizes that are a power of two <= the tile: a row's place inside its page does not depend on
  // the tile, and its page slot is a shift of the row index
  const bool pow2_page = p.page_shift >= 0 && p.page_size <= G::TT;
  // byte strides fit 32 bits (checked by the launcher): one IMAD.WIDE per row instead of 64 x 64
  const uint32_t spage = (uint32_t)(p.kv_stride_page * G::ES), sn = (uint32_t)(p.kv_stride_n * G::ES);
  const int64_t v_minus_k = vb - kb;
  const bool wide = p.page_size * 32 < G::TT;  // a tile can touch more than 32 pages
  // floor(2^32 / page_size): tok / page_size = umulhi(tok, magic) or that + 1 (any tok < 2^32)
  const uint32_t magic = p.page_shift >= 0 ? 0u : (uint32_t)(0x100000000ull / (uint32_t)p.page_size);
  Feeder f;
  feeder_init(f, src);
  int tiles_issued = 0;
  int stage = 0;
  uint32_t phase = 0;
  int cur[R::SLOTS], nxt[R::SLOTS];
  bool have_nxt = false;  // nxt holds the pages of the tile about to be issued (warp-uniform)
  for (int k = 0;; ++k) {
    Segment sg;
    if (feeds) {
      feeder_step(f, src, k, tiles_issued, lane);
      if (!feeder_await(f, src, k, tiles_issued, lane)) break;
      sg = seg_read(src, k);
    } else {
      if (!seg_fetch(src, k, sg)) break;
      __syncwarp();
      if (lane == 0) src.ctl[kQDone1 + (ptid >> 5) - 1 + ((ptid >> 5) > 1 ? 1 : 0)] = k + 1;  // warps 1, 2, 3
    }
    // this thread's chunk column of head sg.kv_head inside a (page, row)
    const char* kcol = kb + ((int64_t)(sg.kv_head / p.kv_head_div) * p.kv_stride_h + c * (16 / G::ES)) * G::ES;
    for (int t = sg.tile_begin; t < sg.tile_end; ++t) {
      if (feeds) feeder_step(f, src, k, tiles_issued, lane);
      if (have_nxt) {
#pragma unroll
        for (int i = 0; i < R::SLOTS; ++i) cur[i] = nxt[i];
      } else {
        cpasync_tile_pages<KVF, HD>(cur, p, sg, t, lane);
      }
      // the next tile's pages: same segment, or the first tile of the next segment that has tiles
      // when it is already published (a non-blocking look: the feeder must not wait here)
      have_nxt = false;
      if (t + 1 < sg.tile_end) {
        cpasync_tile_pages<KVF, HD>(nxt, p, sg, t + 1, lane);
        have_nxt = true;
      } else {
        const bool pub = feeds ? (k + 1 < f.pub) : seg_peek(src, k + 1);
        if (__shfl_sync(0xffffffffu, pub ? 1 : 0, 0) != 0) {
          const Segment sn2 = seg_read(src, k + 1);
          if (sn2.tile_end > sn2.tile_begin) {
            cpasync_tile_pages<KVF, HD>(nxt, p, sn2, sn2.tile_begin, lane);
            have_nxt = true;
          }
        }
      }
      mbar_wait(empty_bar + 8 * stage, phase ^ 1);
      const uint32_t sk = ring + stage * G::STAGE_BYTES;
      const int tok0 = t * G::TT;
      const int nvalid = sg.tok_hi - tok0;  // rows of this tile that hold tokens (>= TT: all)
      int pit0 = 0;
      if (!pow2_page) {
        if (p.page_shift >= 0) pit0 = tok0 >> p.page_shift;
        else { pit0 = (int)__umulhi((uint32_t)tok0, magic); pit0 += (tok0 - pit0 * p.page_size >= p.page_size); }
      }
      auto issue = [&](auto full_tile, auto wide_tile) {
#pragma unroll
        for (int ps = 0; ps < R::PASSES; ++ps) {
          const int row = ps * R::RPP + r0;
          int slot, rip;
          if (pow2_page) { slot = row >> p.page_shift; rip = row & p.page_mask; }
          else {
            const int tok = tok0 + row;
            int pit;
            if (p.page_shift >= 0) pit = tok >> p.page_shift;
            else { pit = (int)__umulhi((uint32_t)tok, magic); pit += (tok - pit * p.page_size >= p.page_size); }
            slot = pit - pit0;
            rip = tok - pit * p.page_size;
          }
          // every lane takes part in the shuffle; rows past the end read a valid lane's (unused) value
          int page = __shfl_sync(0xffffffffu, cur[0], slot & 31);
          if constexpr (R::SLOTS > 1 && decltype(wide_tile)::value) {
            // only tiles that touch more than 32 pages (page size 1): compiled as its own copy of the loop,
            // a warp-uniform branch per row in here cost page size 1 ten per cent
#pragma unroll
            for (int i = 1; i < R::SLOTS; ++i) {
              const int other = __shfl_sync(0xffffffffu, cur[i], slot & 31);
              page = (slot >> 5) == i ? other : page;
            }
          }
          const char* src_k = kcol + ((uint64_t)(uint32_t)page * spage + (uint64_t)(uint32_t)rip * sn);
          const uint32_t d = sk + StageRow<G::TT>(row, ~(G::TT - 1)).at(c);
          if constexpr (decltype(full_tile)::value) {
            cp_async_16_full(d, src_k);
            cp_async_16_full(d + G::PANEL_BYTES, src_k + v_minus_k);
          } else {
            const bool valid = row < nvalid;
            cp_async_16(d, valid ? src_k : kb, valid ? 16u : 0u);
            cp_async_16(d + G::PANEL_BYTES, valid ? src_k + v_minus_k : vb, valid ? 16u : 0u);
          }
        }
      };
      if (wide) {
        if (nvalid >= G::TT) issue(std::true_type{}, std::true_type{});
        else issue(std::false_type{}, std::true_type{});
      } else {
        if (nvalid >= G::TT) issue(std::true_type{}, std::false_type{});
        else issue(std::false_type{}, std::false_type{});
      }
      cp_async_mbar_arrive_noinc(full_bar + 8 * stage);
      if (++stage == p.num_stages) { stage = 0; phase ^= 1; }
      ++tiles_issued;
    }
  }
}

// ---- consumer helpers: 16-bit KV ----------------------------------------------
// Query fragments (B operand of S^T = K Q^T): lane (g = lane/4, t = lane%4) holds head g;
// k-step ks covers d in [16ks, 16ks+16): b0 = q[16ks+2t..+1], b1 = q[16ks+8+2t..+1].
template <int HD>
__device__ __forceinline__ void load_q_frags16(uint32_t (&qf)[HD / 16][2], const void* qrow,
                                               bool head_valid, int t) {
  constexpr int KS = HD / 16;
  if (!head_valid) {
#pragma unroll
    for (int i = 0; i < KS; ++i) qf[i][0] = qf[i][1] = 0u;
    return;
  }
  const uint32_t* q32 = static_cast<const uint32_t*>(qrow);
#pragma unroll
  for (int ks = 0; ks < KS; ++ks) {
    qf[ks][0] = __ldg(q32 + 8 * ks + t);
    qf[ks][1] = __ldg(q32 + 8 * ks + 4 + t);
  }
}

// S^T(16 tokens x 8 heads) for this warp's token slice (rows row0 .. row0+15 of the stage)
template <int QT, int HD>
__device__ __forceinline__ void qk_tile16(float (&s)[4], uint32_t kbase, int row0, int lane,
                                          const uint32_t (&qf)[HD / 16][2], int rmask) {
  constexpr bool BF = (QT == 1);
  constexpr int KS = HD / 16;
  // ldmatrix.x4 address pattern: row = (lane&7) + 8*((lane>>3)&1), chunk += lane>>4
  const int kch = lane >> 4;
  const StageRow<64> kr(row0 + (lane & 7) + 8 * ((lane >> 3) & 1), rmask);
  float s1[4] = {0.f, 0.f, 0.f, 0.f};
  s[0] = s[1] = s[2] = s[3] = 0.f;
#pragma unroll
  for (int ks = 0; ks < KS; ++ks) {
    uint32_t a[4];
    ldmatrix_x4(a, kbase + kr.at(2 * ks + kch));
    if (ks & 1) mma_16816<BF>(s1, a, qf[ks][0], qf[ks][1]);
    else mma_16816<BF>(s, a, qf[ks][0], qf[ks][1]);
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) s[i] += s1[i];
}

// ---- RoPE on the fly (pos_encoding_mode = ROPE_LLAMA, 16-bit KV) -------------------------------
// The reference rotates q at position kv_len - 1 and every K row at its own position inside the
// decode kernel (include/flashinfer/attention/decode.cuh:445-466,72-76; vec_apply_llama_rope,
// include/flashinfer/pos_enc.cuh:79-100): x'[d] = x[d] cos - x[d + D/2] sin, x'[d + D/2] =
// x[d + D/2] cos + x[d] sin with angle = position * rcp_scale * rcp_theta^(2 d / D), d < D/2.
// In the m16n8k16 fragments both members of a rotation pair sit in the SAME lane and register
// slot, k-steps ks and ks + KS/2 apart (A operand: the K rows; B operand: the query), so K is
// rotated in registers between ldmatrix and the MMA and the operand path stays TMA -> smem ->
// ldmatrix -> mma.  Per (token, frequency) one fast sincos (the reference's own __sincosf under
// -use_fast_math) and four FMAs; freq[4 ks + c] belongs to d = 16 ks + 2 t + (c & 1) + 8 (c >> 1).
// The rotated K rows and query are handed to the MMA as f16 whatever the storage type: rounding
// them back to bf16 (8 significant bits) would move the base-2 lse by more than the 1e-3 gate;
// f16 keeps 11 bits and saturates instead of overflowing.
__device__ __forceinline__ uint32_t pack2_f16_sat(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
template <bool BF>
__device__ __forceinline__ float2 unpack2(uint32_t w) {
  if constexpr (BF) return make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xffff0000u));
  else return __half22float2(*reinterpret_cast<const __half2*>(&w));
}

template <int HD>
__device__ __forceinline__ void rope_freqs(float (&freq)[HD / 8], const DecodeParams& p, int t) {
#pragma unroll
  for (int ks = 0; ks < HD / 32; ++ks)
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int d = 16 * ks + 2 * t + (c & 1) + 8 * (c >> 1);
      freq[4 * ks + c] = p.rope_rcp_scale * powf(p.rope_rcp_theta, (float)(2 * d) / (float)HD);
    }
}

// rotate the query fragments of one head at position `pos` (exact sincosf: once per segment)
template <bool BF, int HD>
__device__ __forceinline__ void rope_rotate_q(uint32_t (&qf)[HD / 16][2], const float (&freq)[HD / 8],
                                              int pos) {
  constexpr int H = HD / 32;
#pragma unroll
  for (int ks = 0; ks < H; ++ks)
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      const float2 x = unpack2<BF>(qf[ks][c]), y = unpack2<BF>(qf[ks + H][c]);
      float s0, c0, s1, c1;
      sincosf((float)pos * freq[4 * ks + 2 * c], &s0, &c0);
      sincosf((float)pos * freq[4 * ks + 2 * c + 1], &s1, &c1);
      qf[ks][c] = pack2_f16_sat(x.x * c0 - y.x * s0, x.y * c1 - y.y * s1);
      qf[ks + H][c] = pack2_f16_sat(y.x * c0 + x.x * s0, y.y * c1 + x.y * s1);
    }
}

// S^T(16 tokens x 8 heads) with the K rows rotated at their positions; tok_g = token of row g
template <int QT, int HD>
__device__ __forceinline__ void qk_tile16_rope(float (&s)[4], uint32_t kbase, int row0, int lane,
                                               const uint32_t (&qf)[HD / 16][2],
                                               const float (&freq)[HD / 8], int tok_g, int rmask) {
  constexpr bool BF = (QT == 1);
  constexpr int H = HD / 32;
  const int kch = lane >> 4;
  const StageRow<64> kr(row0 + (lane & 7) + 8 * ((lane >> 3) & 1), rmask);
  const float pos[2] = {(float)tok_g, (float)(tok_g + 8)};
  float s1[4] = {0.f, 0.f, 0.f, 0.f};
  s[0] = s[1] = s[2] = s[3] = 0.f;
#pragma unroll
  for (int ks = 0; ks < H; ++ks) {
    uint32_t a[4], b[4];
    ldmatrix_x4(a, kbase + kr.at(2 * ks + kch));
    ldmatrix_x4(b, kbase + kr.at(2 * (ks + H) + kch));
#pragma unroll
    for (int r = 0; r < 4; ++r) {  // r & 1: row g / g + 8; r >> 1: columns 2t.. / 2t + 8..
      float s0, c0, sn1, c1;
      __sincosf(pos[r & 1] * freq[4 * ks + 2 * (r >> 1)], &s0, &c0);
      __sincosf(pos[r & 1] * freq[4 * ks + 2 * (r >> 1) + 1], &sn1, &c1);
      const float2 x = unpack2<BF>(a[r]), y = unpack2<BF>(b[r]);
      a[r] = pack2_f16_sat(x.x * c0 - y.x * s0, x.y * c1 - y.y * sn1);
      b[r] = pack2_f16_sat(y.x * c0 + x.x * s0, y.y * c1 + x.y * sn1);
    }
    mma_16816<false>(s, a, qf[ks][0], qf[ks][1]);
    mma_16816<false>(s1, b, qf[ks + H][0], qf[ks + H][1]);
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) s[i] += s1[i];
}

// O^T(HD x 8 heads) += V^T . P^T for this warp's 16 tokens; (pb0, pb1) = B fragments of P^T
template <int QT, int HD>
__device__ __forceinline__ void pv_tile16(float (&o)[HD / 16][4], uint32_t vbase, int row0,
                                          int lane, uint32_t pb0, uint32_t pb1, int rmask) {
  constexpr bool BF = (QT == 1);
  // ldmatrix.x4.trans pattern: row = (lane&7) + 8*(lane>>4), chunk += (lane>>3)&1
  const int vch = (lane >> 3) & 1;
  const StageRow<64> vr(row0 + (lane & 7) + 8 * (lane >> 4), rmask);
#pragma unroll
  for (int mt = 0; mt < HD / 16; ++mt) {
    uint32_t a[4];
    ldmatrix_x4_trans(a, vbase + vr.at(2 * mt + vch));
    mma_16816<BF>(o[mt], a, pb0, pb1);
  }
}

// ---- consumer helpers: fp8 KV (native fp8 MMA, no dequantisation) ---------------
// K and V bytes feed mma.m16n8k32 directly as the A operand; the 16-bit query (and later
// the probabilities) are split into an e4m3 "hi" part and an e4m3 residual "lo" part and
// each product is two MMAs into the same accumulator (~8 significant bits, like bf16).
// Query fragments: lane (g, t) holds head g; k-block c covers d in [32c, 32c+32):
//   b0 = q[32c+4t .. +3], b1 = q[32c+16+4t .. +3] (4 e4m3 bytes each).
// The query row is scaled so that its amax maps to 256; `dq` is the inverse scale.
template <int QT, int HD>
__device__ __forceinline__ void load_q_frags8(uint32_t (&qhi)[HD / 32][2],
                                              uint32_t (&qlo)[HD / 32][2],
                                              uint32_t (&ql2)[HD / 32][2], float& dq,
                                              const void* qrow, bool head_valid, int t) {
  constexpr int KB = HD / 32;
  float x[KB][2][4];
  float amax = 0.f;
  const uint2* q64 = static_cast<const uint2*>(qrow);
#pragma unroll
  for (int c = 0; c < KB; ++c)
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      uint2 v = make_uint2(0u, 0u);
      if (head_valid) v = __ldg(q64 + 8 * c + 4 * h + t);
      if constexpr (QT == 1) {
        x[c][h][0] = __uint_as_float(v.x << 16); x[c][h][1] = __uint_as_float(v.x & 0xffff0000u);
        x[c][h][2] = __uint_as_float(v.y << 16); x[c][h][3] = __uint_as_float(v.y & 0xffff0000u);
      } else {
        const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&v.x));
        const float2 b = __half22float2(*reinterpret_cast<const __half2*>(&v.y));
        x[c][h][0] = a.x; x[c][h][1] = a.y; x[c][h][2] = b.x; x[c][h][3] = b.y;
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) amax = fmaxf(amax, fabsf(x[c][h][i]));
    }
  amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
  amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
  const bool nz = amax > 0.f && amax < 3.0e38f;
  const float sq = nz ? 256.f / amax : 1.f;
  dq = nz ? amax * (1.f / 256.f) : 1.f;
#pragma unroll
  for (int c = 0; c < KB; ++c)
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      // three e4m3 terms (~12 significant bits): the base-2 LSE must hold to 1e-3
      uint32_t t01[3], t23[3];
      split3_e4m3x2(t01, x[c][h][0] * sq, x[c][h][1] * sq);
      split3_e4m3x2(t23, x[c][h][2] * sq, x[c][h][3] * sq);
      qhi[c][h] = t01[0] | (t23[0] << 16);
      qlo[c][h] = t01[1] | (t23[1] << 16);
      ql2[c][h] = t01[2] | (t23[2] << 16);
    }
}

// S^T for this warp's 32 tokens (two 16-token M tiles) at stage rows row0 .. row0+31
template <int KVF, int HD>
__device__ __forceinline__ void qk_tile8(float (&s)[2][4], uint32_t kbase, int row0, int lane,
                                         const uint32_t (&qhi)[HD / 32][2],
                                         const uint32_t (&qlo)[HD / 32][2],
                                         const uint32_t (&ql2)[HD / 32][2], int rmask) {
  constexpr int KB = HD / 32;
  const int krow = row0 + (lane & 7) + 8 * ((lane >> 3) & 1), kch = lane >> 4;
  const StageRow<128> kr0(krow, rmask), kr1(krow + 16, rmask);
  float s2[2][4];
#pragma unroll
  for (int mb = 0; mb < 2; ++mb)
#pragma unroll
    for (int i = 0; i < 4; ++i) s[mb][i] = s2[mb][i] = 0.f;
#pragma unroll
  for (int c = 0; c < KB; ++c)
#pragma unroll
    for (int mb = 0; mb < 2; ++mb) {
      uint32_t a[4];
      ldmatrix_x4(a, kbase + (mb ? kr1 : kr0).at(2 * c + kch));
      mma_16832_fp8<KVF == 2>(s[mb], a, qhi[c][0], qhi[c][1]);
      mma_16832_fp8<KVF == 2>(s2[mb], a, qlo[c][0], qlo[c][1]);
      mma_16832_fp8<KVF == 2>(s2[mb], a, ql2[c][0], ql2[c][1]);
    }
#pragma unroll
  for (int mb = 0; mb < 2; ++mb)
#pragma unroll
    for (int i = 0; i < 4; ++i) s[mb][i] += s2[mb][i];
}

// O^T += V^T . P^T over this warp's 32 tokens.  ldmatrix.trans hands out byte pairs
// [tok 2t: d 2g, d 2g+1 | tok 2t+1: d 2g, d 2g+1] per 8-token group; prmt regroups them by d so
// that MMA row g <-> d = 16c + 2g, row g+8 <-> d = 16c + 2g + 1 and the k slots 4t..4t+3 of
// each half <-> tokens {2t, 2t+1, 8+2t, 9+2t} (+16).  bh/bl: hi / lo B fragments of P^T in the
// same token order.
template <int KVF, int HD>
__device__ __forceinline__ void pv_tile8(float (&o)[HD / 16][4], uint32_t vbase, int row0, int lane,
                                         const uint32_t (&bh)[2], const uint32_t (&bl)[2], int rmask) {
  const StageRow<128> vr(row0 + lane, rmask);  // matrix j = lane>>3 covers tokens 8j .. 8j+7
#pragma unroll
  for (int c = 0; c < HD / 16; ++c) {
    uint32_t r[4], a[4];
    ldmatrix_x4_trans(r, vbase + vr.at(c));
    a[0] = prmt(r[0], r[1], 0x6420u);
    a[1] = prmt(r[0], r[1], 0x7531u);
    a[2] = prmt(r[2], r[3], 0x6420u);
    a[3] = prmt(r[2], r[3], 0x7531u);
    mma_16832_fp8<KVF == 2>(o[c], a, bh[0], bh[1]);
    mma_16832_fp8<KVF == 2>(o[c], a, bl[0], bl[1]);
  }
}

// head-dim index of accumulator row (mt, hi) for lane group g
template <int KVF>
__device__ __forceinline__ int o_dim(int mt, int hi, int g) {
  if constexpr (KVF == 0) return 16 * mt + g + 8 * hi;
  else return 16 * mt + 2 * g + hi;
}

// ---- fused split-KV merge ---------------------------------------------------------
// Run by ONE warp of the CTA whose partial arrived last at a split unit: combines the unit's
// `count` normalised f32 partials (written by this and other CTAs, hence .cg loads) into the
// final o / lse.  Same operator as the reference's merge (cascade.cuh:366-467, state.cuh:53-62).
__device__ __forceinline__ float ldcg_f32(const float* p) {
  float v;
  asm volatile("ld.global.cg.f32 %0, [%1];" : "=f"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ float4 ldcg_f32x4(const float* p) {
  float4 v;
  asm volatile("ld.global.cg.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
// ---- multi-row units (DecodeParams::q_rows = n > 1, run(q_len_per_req = n)) -------------------
// A request's n new query tokens share its KV stream: the unit's GS query "heads" are (row, head)
// pairs.  Column c of virtual KV head vh is column cc = (vh % div) * GS + c of the real KV head
// kv = vh / div, i.e. row cc / G of the n rows and query head kv * G + cc % G (G = q_group, the
// model's GQA group).  Row r is query token r of the request and attends all but the last
// n - 1 - r tokens of the cache (causal among the new tokens).
struct RowHead {
  int row, head;  // row of the request's n query rows; real query head
  __device__ __forceinline__ int64_t token(const DecodeParams& p, int request) const {
    return (int64_t)request * p.q_rows + row;  // row of q / o / lse
  }
};
__device__ __forceinline__ RowHead row_head(const DecodeParams& p, int kv_head, int c) {
  const int cc = (kv_head % p.kv_head_div) * p.group_size + c;
  RowHead r;
  r.row = cc / p.q_group;
  r.head = (kv_head / p.kv_head_div) * p.q_group + (cc - r.row * p.q_group);
  return r;
}

// ctid: 0 .. 127 (the four consumer warps).  One item = (head, 4 dims); loads are issued in
// batches of four independent .cg reads so that the unit costs ~two L2 round trips.
template <int QT, int HD, bool MROWS = false>
__device__ __forceinline__ void merge_unit_cta(const DecodeParams& p, int request, int kv_head,
                                               int slot_begin, int count, int ctid) {
  const int GS = p.group_size;
  constexpr int CPH = HD / 4;  // float4 chunks per head row
  // One pass with a running maximum (rescaled on the fly) over batches of four partials: the lse
  // and o loads of a batch are issued together, so a unit of <= 4 partials costs ONE L2 round trip.
#pragma unroll 1
  for (int item = ctid; item < GS * CPH; item += kConsumerWarps * 32) {
    const int h = item / CPH, c = item - h * CPH;
    const float* lse = p.tmp_lse + (int64_t)slot_begin * GS + h;
    const float* row = p.tmp_o + ((int64_t)slot_begin * GS + h) * HD + 4 * c;
    float mx = kNegInf, d = 0.f;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 1
    for (int k0 = 0; k0 < count; k0 += 4) {
      float s4[4];
      float4 x4[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const bool ok = k0 + j < count;
        s4[j] = ok ? __ldcg(lse + (int64_t)(k0 + j) * GS) : kNegInf;
        x4[j] = ok ? __ldcg(reinterpret_cast<const float4*>(row + (int64_t)(k0 + j) * GS * HD))
                   : make_float4(0.f, 0.f, 0.f, 0.f);
      }
      float mn = mx;
#pragma unroll
      for (int j = 0; j < 4; ++j) mn = fmaxf(mn, s4[j]);
      const float sc = ex2f(mx - mn);  // 1 when the maximum did not move, 0 out of the sentinel
      mx = mn;
      d *= sc;
      acc.x *= sc; acc.y *= sc; acc.z *= sc; acc.w *= sc;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float w = k0 + j < count ? ex2f(s4[j] - mn) : 0.f;
        d += w;
        acc.x += w * x4[j].x; acc.y += w * x4[j].y; acc.z += w * x4[j].z; acc.w += w * x4[j].w;
      }
    }
    int64_t hq = (int64_t)request * p.num_qo_heads + kv_head * GS + h;
    int sink_head = kv_head * GS + h;
    if constexpr (MROWS) {
      const RowHead rh = row_head(p, kv_head, h);
      sink_head = rh.head;
      hq = rh.token(p, request) * (p.num_qo_heads / p.q_rows) + rh.head;
    }
    if (p.sinks != nullptr && mx > kNegInf) d += ex2f(__ldg(p.sinks + sink_head) * 1.4426950408889634f - mx);
    const float inv = p.v_scale / d;
    uint2 u;
    u.x = pack2<QT == 1>(acc.x * inv, acc.y * inv);
    u.y = pack2<QT == 1>(acc.z * inv, acc.w * inv);
    *reinterpret_cast<uint2*>(static_cast<uint16_t*>(p.o) + hq * HD + 4 * c) = u;
    // all partials empty: weights are 1 and o stays 0; clamp lse to the sentinel
    if (c == 0 && p.lse != nullptr) p.lse[hq] = mx > kNegInf ? mx + lg2f(d) : kNegInf;
  }
}

// Deferred arrival check.  A partial's arrival is an acq_rel atomic on the unit's counter whose
// RESULT is not looked at until the CTA's next segment epilogue (or the end of its sequence), so
// the L2 round trip hides behind the next segment.  The CTA whose arrival turns out to have been
// the last one re-arms the counter and merges the unit; all 128 consumer threads take part, so
// thread 0 posts the decision in shared memory ahead of a CTA barrier.
struct PendingMerge {
  int mu, old, cnt, slot, req, head;  // meaningful in thread 0 only
};
__device__ __forceinline__ void merge_arrive(const DecodeParams& p, PendingMerge& pm, const Segment& sg) {
  pm.old = atom_add_acq_rel_gpu(p.merge_counters + sg.merge_unit, 1);
  pm.mu = sg.merge_unit; pm.cnt = sg.merge_count; pm.slot = sg.merge_slot;
  pm.req = sg.request; pm.head = sg.kv_head;
}
// thread 0, before a barrier of the consumer threads
__device__ __forceinline__ void merge_post(const DecodeParams& p, PendingMerge& pm, float* scratch) {
  volatile int* d = reinterpret_cast<volatile int*>(scratch + kMergeDescFloat);
  const bool go = pm.mu >= 0 && pm.old == pm.cnt - 1;
  if (go) {
    p.merge_counters[pm.mu] = 0;
    d[1] = pm.req; d[2] = pm.head; d[3] = pm.slot; d[4] = pm.cnt;
  }
  d[0] = go ? 1 : 0;
  pm.mu = -1;
}
// all consumer threads, after that barrier (and before the next one)
template <int QT, int HD, bool MROWS = false>
__device__ __forceinline__ void merge_run_posted(const DecodeParams& p, const float* scratch, int ctid) {
  const volatile int* d = reinterpret_cast<const volatile int*>(scratch + kMergeDescFloat);
  if (d[0]) {
    __threadfence();
    merge_unit_cta<QT, HD, MROWS>(p, d[1], d[2], d[3], d[4], ctid);
    FI_TRACE_IF(ctid == 0, 9);
  }
}

// ---- consumer: 4 warps, online softmax + MMA ---------------------------------
// Per-warp running state per query head j (two heads per lane: 2t, 2t+1):
//   m[j]  reference exponent of o / l (16-bit KV: the running max; fp8 KV: max(block max,
//         running max - 20), so that probabilities quantised to e4m3 are scaled by their own
//         32-token block's maximum and nothing within 2^-18 of it is flushed),
//   l[j]  sum of exp2(x - m[j]),  o[..] = sum of exp2(x - m[j]) v  (x 256 for fp8 KV).
template <int QT, int KVF, int HD, int LOADER, bool ROPE = false, bool MROWS = false>
__device__ __forceinline__ void consume(const DecodeParams& p, const SegSource& src, uint32_t ring,
                                        uint32_t full_bar, uint32_t empty_bar, float* scratch_o,
                                        float* scratch_ml) {
  using G = TileGeom<KVF, HD>;
  constexpr bool BF = (QT == 1);  // operand type of P (16-bit KV path)
  constexpr int MT = HD / 16;
  constexpr int TPW = G::TPW;
  constexpr int NQ = (KVF == 0) ? HD / 16 : HD / 32;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t = lane & 3;
  const int GS = p.group_size;
  const bool use_cap = p.soft_cap > 0.f;
  const float cap_pre = use_cap ? p.sm_scale / p.soft_cap : 0.f;
  const float cap_post = use_cap ? p.soft_cap * 1.4426950408889634f : 0.f;
  const int row0 = warp * TPW;
  // stage layout (StageRow): token blocks of rpb rows, K rows then V rows; the cp.async loader
  // writes whole-tile blocks
  const int rpb = LOADER == 0 ? p.rows_per_box : G::TT;
  const int rmask = ~(rpb - 1);
  const uint32_t vdelta = (uint32_t)rpb * 128u;
  static_assert(!ROPE || KVF == 0, "in-register RoPE is built for 16-bit KV");
  static_assert(!MROWS || (KVF == 0 && !ROPE), "multi-row units: 16-bit KV, no RoPE");
  [[maybe_unused]] float freq[ROPE ? HD / 8 : 1];
  if constexpr (ROPE) rope_freqs<HD>(freq, p, t);

  int stage = 0;
  uint32_t phase = 0;
  PendingMerge pm;
  pm.mu = -1; pm.old = pm.cnt = pm.slot = pm.req = pm.head = 0;
  // 16-bit KV: the query fragments of segment k + 1 are requested at the start of segment k when
  // that segment is already published (warp-uniform decision), so a boundary costs no round trip
  [[maybe_unused]] uint32_t qn[NQ][2];
  [[maybe_unused]] bool have_qn = false;
  for (int k = 0;; ++k) {
    Segment sg;
    if (!seg_fetch(src, k, sg)) break;
    uint32_t qf[NQ][2];
    uint32_t ql[KVF == 0 ? 1 : NQ][2], ql2[KVF == 0 ? 1 : NQ][2];  // fp8: residual query terms
    float mul[2], pre[2];               // logit scale per head column (2t, 2t+1)
    // multi-row units: tokens at the end of the cache that head column 2t + j does not see
    [[maybe_unused]] int hid[2] = {0, 0};
    {
      const char* qrow = static_cast<const char*>(p.q) +
                         ((int64_t)sg.request * p.q_stride_n +
                          (int64_t)(sg.kv_head * GS + g) * p.q_stride_h) * 2;
      if constexpr (MROWS) {
        const RowHead rh = row_head(p, sg.kv_head, g);
        qrow = static_cast<const char*>(p.q) +
               (rh.token(p, sg.request) * p.q_stride_n + (int64_t)rh.head * p.q_stride_h) * 2;
        if (!p.q_rows_noncausal) {
          hid[0] = p.q_rows - 1 - row_head(p, sg.kv_head, 2 * t).row;
          hid[1] = p.q_rows - 1 - row_head(p, sg.kv_head, 2 * t + 1).row;
        }
      }
      if constexpr (KVF == 0) {
        if (have_qn) {
#pragma unroll
          for (int i = 0; i < NQ; ++i) { qf[i][0] = qn[i][0]; qf[i][1] = qn[i][1]; }
        } else {
          load_q_frags16<HD>(qf, qrow, g < GS, t);
          if constexpr (ROPE) rope_rotate_q<BF, HD>(qf, freq, sg.kv_len - 1);
        }
        have_qn = __shfl_sync(0xffffffffu, seg_peek(src, k + 1) ? 1 : 0, 0) != 0;
        if (have_qn) {
          const Segment sn = seg_read(src, k + 1);
          const char* nrow = static_cast<const char*>(p.q) +
                             ((int64_t)sn.request * p.q_stride_n +
                              (int64_t)(sn.kv_head * GS + g) * p.q_stride_h) * 2;
          if constexpr (MROWS) {
            const RowHead rn = row_head(p, sn.kv_head, g);
            nrow = static_cast<const char*>(p.q) +
                   (rn.token(p, sn.request) * p.q_stride_n + (int64_t)rn.head * p.q_stride_h) * 2;
          }
          load_q_frags16<HD>(qn, nrow, g < GS, t);
          if constexpr (ROPE) rope_rotate_q<BF, HD>(qn, freq, sn.kv_len - 1);
        }
        ql[0][0] = ql[0][1] = ql2[0][0] = ql2[0][1] = 0u;
        mul[0] = mul[1] = p.sm_scale_log2;
        pre[0] = pre[1] = cap_pre;
      } else {
        float dq;
        load_q_frags8<QT, HD>(qf, ql, ql2, dq, qrow, g < GS, t);
        const float d0 = __shfl_sync(0xffffffffu, dq, 8 * t);      // head 2t lives in lanes 8t..8t+3
        const float d1 = __shfl_sync(0xffffffffu, dq, 8 * t + 4);  // head 2t+1
        mul[0] = d0 * p.sm_scale_log2; mul[1] = d1 * p.sm_scale_log2;
        pre[0] = d0 * cap_pre; pre[1] = d1 * cap_pre;
      }
    }
    // ALiBi (reference variants.cuh:68-71, qo_idx = 0 as decode.cuh:95 passes it):
    // logit = q.k * sm_scale + slope[head] * kv position; alc is the slope in the units of the
    // expression it is added to (base-2 logits, or the tanh argument when soft-capped, where the
    // reference multiplies the biased logit by sm_scale / cap once more)
    float alc[2] = {0.f, 0.f};
    if (p.alibi_slopes != nullptr) {
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int h = 2 * t + j;
        const int ah = MROWS ? row_head(p, sg.kv_head, h).head : sg.kv_head * GS + h;
        const float a = h < GS ? __ldg(p.alibi_slopes + ah) : 0.f;
        alc[j] = a * (use_cap ? cap_pre : 1.4426950408889634f);
        pre[j] *= p.sm_scale;
      }
    }
    float o[MT][4];
#pragma unroll
    for (int i = 0; i < MT; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
    float m[2] = {kNegInf, kNegInf}, l[2] = {0.f, 0.f};
    float mrun[2] = {kNegInf, kNegInf};  // fp8 only: true running max

    for (int tile = sg.tile_begin; tile < sg.tile_end; ++tile) {
      mbar_wait(full_bar + 8 * stage, phase);
      FI_TRACE_IF(threadIdx.x == 0 && k == 0 && tile == sg.tile_begin, 5);
      const int tok0 = tile * G::TT + row0;
      const int vcount = min(TPW, sg.tok_hi - tok0);
      if (vcount > 0 && tok0 + TPW > sg.tok_lo) {  // warp-uniform
        const uint32_t kbase = ring + stage * G::STAGE_BYTES;
        const uint32_t vbase = kbase + vdelta;
        if constexpr (LOADER == 0) {
          if (vcount < TPW) {  // rows the TMA did not (validly) fill: zero V so 0 * junk != NaN
            for (int i = lane; i < (TPW - vcount) * G::CHUNKS; i += 32) {
              const int row = row0 + vcount + i / G::CHUNKS;
              const uint32_t a = vbase + StageRow<G::TT>(row, rmask).at(i % G::CHUNKS);
              asm volatile("st.shared.v4.b32 [%0], {%1,%1,%1,%1};" ::"r"(a), "r"(0u) : "memory");
            }
            __syncwarp();
          }
        }
        if constexpr (KVF == 0) {
          float s[4];
          if constexpr (ROPE) qk_tile16_rope<QT, HD>(s, kbase, row0, lane, qf, freq, tok0 + g, rmask);
          else qk_tile16<QT, HD>(s, kbase, row0, lane, qf, rmask);
          // s[0],s[1]: token g, heads 2t,2t+1;  s[2],s[3]: token g+8
          const int ta = tok0 + g, tb = ta + 8;
          const bool va = ta >= sg.tok_lo && ta < sg.tok_hi;
          const bool vb = tb >= sg.tok_lo && tb < sg.tok_hi;
          const float pf[2] = {(float)ta, (float)tb};
          float x[4];
#pragma unroll
          for (int i = 0; i < 4; ++i)
            x[i] = use_cap ? cap_post * tanhf_approx(s[i] * pre[i & 1] + alc[i & 1] * pf[i >> 1])
                           : s[i] * mul[i & 1] + alc[i & 1] * pf[i >> 1];
          [[maybe_unused]] bool vc[4] = {va, va, vb, vb};  // per (token, head column)
          if constexpr (MROWS) {
            vc[0] = va && ta < sg.tok_hi - hid[0]; vc[1] = va && ta < sg.tok_hi - hid[1];
            vc[2] = vb && tb < sg.tok_hi - hid[0]; vc[3] = vb && tb < sg.tok_hi - hid[1];
            x[0] = vc[0] ? x[0] : kNegInf; x[1] = vc[1] ? x[1] : kNegInf;
            x[2] = vc[2] ? x[2] : kNegInf; x[3] = vc[3] ? x[3] : kNegInf;
          } else {
          x[0] = va ? x[0] : kNegInf; x[1] = va ? x[1] : kNegInf;
          x[2] = vb ? x[2] : kNegInf; x[3] = vb ? x[3] : kNegInf;
          }
          float mx[2] = {fmaxf(x[0], x[2]), fmaxf(x[1], x[3])};
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            mx[j] = fmaxf(mx[j], __shfl_xor_sync(0xffffffffu, mx[j], 4));
            mx[j] = fmaxf(mx[j], __shfl_xor_sync(0xffffffffu, mx[j], 8));
            mx[j] = fmaxf(mx[j], __shfl_xor_sync(0xffffffffu, mx[j], 16));
          }
          float sc[2];
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            const float mn = fmaxf(m[j], mx[j]);
            sc[j] = ex2f(m[j] - mn);
            m[j] = mn;
          }
          float pr[4];
          if constexpr (MROWS) {
            pr[0] = vc[0] ? ex2f(x[0] - m[0]) : 0.f;
            pr[1] = vc[1] ? ex2f(x[1] - m[1]) : 0.f;
            pr[2] = vc[2] ? ex2f(x[2] - m[0]) : 0.f;
            pr[3] = vc[3] ? ex2f(x[3] - m[1]) : 0.f;
          } else {
          pr[0] = va ? ex2f(x[0] - m[0]) : 0.f;
          pr[1] = va ? ex2f(x[1] - m[1]) : 0.f;
          pr[2] = vb ? ex2f(x[2] - m[0]) : 0.f;
          pr[3] = vb ? ex2f(x[3] - m[1]) : 0.f;
          }
          l[0] = l[0] * sc[0] + pr[0] + pr[2];
          l[1] = l[1] * sc[1] + pr[1] + pr[3];
#pragma unroll
          for (int i = 0; i < MT; ++i) {
            o[i][0] *= sc[0]; o[i][1] *= sc[1]; o[i][2] *= sc[0]; o[i][3] *= sc[1];
          }
          const uint32_t pb0 = movmatrix_trans(pack2<BF>(pr[0], pr[1]));
          const uint32_t pb1 = movmatrix_trans(pack2<BF>(pr[2], pr[3]));
          pv_tile16<QT, HD>(o, vbase, row0, lane, pb0, pb1, rmask);
        } else {
          float s[2][4];
          qk_tile8<KVF, HD>(s, kbase, row0, lane, qf, ql, ql2, rmask);
          // s[mb][0..1]: token 16mb+g, heads 2t,2t+1;  s[mb][2..3]: token 16mb+g+8
          bool valid[2][2];
          float mx[2] = {kNegInf, kNegInf};
#pragma unroll
          for (int mb = 0; mb < 2; ++mb)
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              const int tk = tok0 + 16 * mb + 8 * h + g;
              valid[mb][h] = tk >= sg.tok_lo && tk < sg.tok_hi;
#pragma unroll
              for (int j = 0; j < 2; ++j) {
                float& x = s[mb][2 * h + j];
                x = use_cap ? cap_post * tanhf_approx(x * pre[j] + alc[j] * (float)tk)
                            : x * mul[j] + alc[j] * (float)tk;
                x = valid[mb][h] ? x : kNegInf;
                mx[j] = fmaxf(mx[j], x);
              }
            }
          float sc[2];
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            mx[j] = fmaxf(mx[j], __shfl_xor_sync(0xffffffffu, mx[j], 4));
            mx[j] = fmaxf(mx[j], __shfl_xor_sync(0xffffffffu, mx[j], 8));
            mx[j] = fmaxf(mx[j], __shfl_xor_sync(0xffffffffu, mx[j], 16));
            mrun[j] = fmaxf(mrun[j], mx[j]);
            const float mn = fmaxf(mx[j], mrun[j] - 20.f);
            sc[j] = ex2f(m[j] - mn);  // <= 2^20
            m[j] = mn;
          }
          uint32_t tq[4];  // tq[2mb+h]: transposed (hi, lo) byte pairs of token group 8(2mb+h)..+7
          float ls[2] = {0.f, 0.f};
#pragma unroll
          for (int mb = 0; mb < 2; ++mb)
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              const float p0 = valid[mb][h] ? ex2f(s[mb][2 * h] - m[0]) : 0.f;
              const float p1 = valid[mb][h] ? ex2f(s[mb][2 * h + 1] - m[1]) : 0.f;
              ls[0] += p0; ls[1] += p1;
              // [hi(h0), hi(h1), lo(h0), lo(h1)] -> [hi(h0), lo(h0), hi(h1), lo(h1)]
              const uint32_t w = prmt(split_e4m3x2(p0 * 256.f, p1 * 256.f), 0u, 0x3120u);
              tq[2 * mb + h] = movmatrix_trans(w);
            }
          l[0] = l[0] * sc[0] + ls[0];
          l[1] = l[1] * sc[1] + ls[1];
#pragma unroll
          for (int i = 0; i < MT; ++i) {
            o[i][0] *= sc[0]; o[i][1] *= sc[1]; o[i][2] *= sc[0]; o[i][3] *= sc[1];
          }
          // after the transpose lane (g, t) holds, for head g, [tok 2t: hi, lo | tok 2t+1: hi, lo]
          const uint32_t bh[2] = {prmt(tq[0], tq[1], 0x6420u), prmt(tq[2], tq[3], 0x6420u)};
          const uint32_t bl[2] = {prmt(tq[0], tq[1], 0x7531u), prmt(tq[2], tq[3], 0x7531u)};
          pv_tile8<KVF, HD>(o, vbase, row0, lane, bh, bl, rmask);
        }
        if constexpr (LOADER == 0) {
          if (vcount < TPW) fence_proxy_async_smem();  // generic zero-fill before next TMA write
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(empty_bar + 8 * stage);
      if (++stage == p.num_stages) { stage = 0; phase ^= 1; }
    }

    FI_TRACE_IF(threadIdx.x == 0, 6);
    // ---- segment epilogue: merge the 4 warp states, write o/lse or a partial ----
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      l[j] += __shfl_xor_sync(0xffffffffu, l[j], 4);
      l[j] += __shfl_xor_sync(0xffffffffu, l[j], 8);
      l[j] += __shfl_xor_sync(0xffffffffu, l[j], 16);
    }
    if (g == 0) {
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        scratch_ml[(warp * kMaxGroup + 2 * t + j) * 2 + 0] = m[j];
        scratch_ml[(warp * kMaxGroup + 2 * t + j) * 2 + 1] = l[j];
      }
    }
    if (threadIdx.x == 0) merge_post(p, pm, scratch_ml);  // arrival of the previous partial
    named_bar_sync(1, kConsumerWarps * 32);
    if (threadIdx.x == 0) src.ctl[kQDone0] = k + 1;  // every warp holds its copy of segment k
    FI_TRACE_IF(threadIdx.x == 0, 10);
    merge_run_posted<QT, HD, MROWS>(p, scratch_ml, threadIdx.x);
    float M[2], L[2], f[2];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      float mm = kNegInf;
#pragma unroll
      for (int w = 0; w < kConsumerWarps; ++w)
        mm = fmaxf(mm, scratch_ml[(w * kMaxGroup + 2 * t + j) * 2]);
      float ll = 0.f;
#pragma unroll
      for (int w = 0; w < kConsumerWarps; ++w)
        ll += scratch_ml[(w * kMaxGroup + 2 * t + j) * 2 + 1] *
              ex2f(scratch_ml[(w * kMaxGroup + 2 * t + j) * 2] - mm);
      M[j] = mm; L[j] = ll; f[j] = ex2f(m[j] - mm);
    }
#pragma unroll
    for (int i = 0; i < MT; ++i) {
      o[i][0] *= f[0]; o[i][1] *= f[1]; o[i][2] *= f[0]; o[i][3] *= f[1];
    }
    if (warp != 0) {
      float* dst = scratch_o + (warp - 1) * (MT * 4 * 32) + lane;
#pragma unroll
      for (int i = 0; i < MT; ++i)
#pragma unroll
        for (int r = 0; r < 4; ++r) dst[(i * 4 + r) * 32] = o[i][r];
    }
    named_bar_sync(2, kConsumerWarps * 32);
    FI_TRACE_IF(threadIdx.x == 0, 11);
    if (warp == 0) {
#pragma unroll
      for (int w = 0; w < kConsumerWarps - 1; ++w) {
        const float* s2 = scratch_o + w * (MT * 4 * 32) + lane;
#pragma unroll
        for (int i = 0; i < MT; ++i)
#pragma unroll
          for (int r = 0; r < 4; ++r) o[i][r] += s2[(i * 4 + r) * 32];
      }
      const bool direct = sg.slot < 0;
      float inv[2], lse2[2];
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const bool any = L[j] > 0.f;
        if (direct && any && p.sinks != nullptr && 2 * t + j < GS) {
          const int sh = MROWS ? row_head(p, sg.kv_head, 2 * t + j).head : sg.kv_head * GS + 2 * t + j;
          L[j] += ex2f(__ldg(p.sinks + sh) * 1.4426950408889634f - M[j]);
        }
        // fp8 KV: the accumulator carries the x256 scale of the quantised probabilities
        inv[j] = any ? (direct ? p.v_scale : 1.f) * (KVF == 0 ? 1.f : 1.f / 256.f) / L[j] : 0.f;
        lse2[j] = any ? M[j] + lg2f(L[j]) : kNegInf;
      }
      const int h0 = 2 * t;
      if (direct) {
        const int64_t hq = (int64_t)sg.request * p.num_qo_heads + sg.kv_head * GS;
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          if (h0 + j < GS) {
            int64_t oh = hq + h0 + j;  // row of o / lse: (request, query head)
            if constexpr (MROWS) {
              const RowHead rh = row_head(p, sg.kv_head, h0 + j);
              oh = rh.token(p, sg.request) * (p.num_qo_heads / p.q_rows) + rh.head;
            }
            if (QT == 1) {
              __nv_bfloat16* orow = static_cast<__nv_bfloat16*>(p.o) + oh * HD;
#pragma unroll
              for (int i = 0; i < MT; ++i) {
                orow[o_dim<KVF>(i, 0, g)] = __float2bfloat16_rn(o[i][j] * inv[j]);
                orow[o_dim<KVF>(i, 1, g)] = __float2bfloat16_rn(o[i][2 + j] * inv[j]);
              }
            } else {
              __half* orow = static_cast<__half*>(p.o) + oh * HD;
#pragma unroll
              for (int i = 0; i < MT; ++i) {
                orow[o_dim<KVF>(i, 0, g)] = __float2half_rn(o[i][j] * inv[j]);
                orow[o_dim<KVF>(i, 1, g)] = __float2half_rn(o[i][2 + j] * inv[j]);
              }
            }
            if (p.lse != nullptr && g == 0) p.lse[oh] = lse2[j];
          }
        }
      } else {
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          if (h0 + j < GS) {
            float* orow = p.tmp_o + ((int64_t)sg.slot * GS + h0 + j) * HD;
#pragma unroll
            for (int i = 0; i < MT; ++i) {
              orow[o_dim<KVF>(i, 0, g)] = o[i][j] * inv[j];
              orow[o_dim<KVF>(i, 1, g)] = o[i][2 + j] * inv[j];
            }
            if (g == 0) p.tmp_lse[(int64_t)sg.slot * GS + h0 + j] = lse2[j];
          }
        }
        if (sg.merge_unit >= 0) {  // publish the partial: count the arrival, look at it later
          __syncwarp();
          if (lane == 0) merge_arrive(p, pm, sg);
        }
      }
    }
    FI_TRACE_IF(threadIdx.x == 0, 7);
  }
  // end of this CTA's sequence: the last arrival is still unchecked
  if (threadIdx.x == 0) merge_post(p, pm, scratch_ml);
  named_bar_sync(1, kConsumerWarps * 32);
  merge_run_posted<QT, HD, MROWS>(p, scratch_ml, threadIdx.x);
  FI_TRACE_IF(threadIdx.x == 0, 8);
}

// ---- consumer for multi-row units of up to 16 (row, head) columns: TWO N tiles per warp ----------
// (FI_PLAN_WIDE_UNITS plans, 16-bit KV, TMA loader.)  The warp's 16 tokens meet both 8-column tiles per
// ldmatrix: one K fragment feeds two Q.K MMAs, one V fragment two P.V MMAs, so 16 columns cost the KV
// stream once.  A separate function on purpose: the single-tile consumer above is the headline kernel
// and must not move by a register.  Structure and epilogue as in consume(); no query prefetch across
// segments (its fragments would hold 32 more registers).
template <int QT, int HD, int LOADER>
__device__ __forceinline__ void consume_rows2(const DecodeParams& p, const SegSource& src, uint32_t ring,
                                              uint32_t full_bar, uint32_t empty_bar, float* scratch_o,
                                              float* scratch_ml) {
  using G = TileGeom<0, HD>;
  constexpr bool BF = (QT == 1);
  constexpr int MT = HD / 16, KS = HD / 16, TPW = G::TPW, NT = 2;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t = lane & 3;
  const int GS = p.group_size;  // 9 .. 16 columns
  const bool use_cap = p.soft_cap > 0.f;
  const float cap_pre = use_cap ? p.sm_scale / p.soft_cap : 0.f;
  const float cap_post = use_cap ? p.soft_cap * 1.4426950408889634f : 0.f;
  const int row0 = warp * TPW;
  const int rpb = LOADER == 0 ? p.rows_per_box : G::TT;
  const int rmask = ~(rpb - 1);
  const uint32_t vdelta = (uint32_t)rpb * 128u;
  const int hq_real = p.num_qo_heads / p.q_rows;

  int stage = 0;
  uint32_t phase = 0;
  PendingMerge pm;
  pm.mu = -1; pm.old = pm.cnt = pm.slot = pm.req = pm.head = 0;
  for (int k = 0;; ++k) {
    Segment sg;
    if (!seg_fetch(src, k, sg)) break;
    uint32_t qf[NT][KS][2];
    int hid[NT][2];
    float alc[NT][2];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const int col = 8 * nt + g;
      const RowHead rh = row_head(p, sg.kv_head, col);
      const char* qrow = static_cast<const char*>(p.q) +
                         (rh.token(p, sg.request) * p.q_stride_n + (int64_t)rh.head * p.q_stride_h) * 2;
      load_q_frags16<HD>(qf[nt], qrow, col < GS, t);
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int c = 8 * nt + 2 * t + j;
        const RowHead rc = row_head(p, sg.kv_head, c);
        hid[nt][j] = p.q_rows_noncausal ? 0 : p.q_rows - 1 - rc.row;
        alc[nt][j] = 0.f;
        if (p.alibi_slopes != nullptr && c < GS)
          alc[nt][j] = __ldg(p.alibi_slopes + rc.head) * (use_cap ? cap_pre : 1.4426950408889634f);
      }
    }
    const float mul = p.sm_scale_log2;
    const float pre = p.alibi_slopes != nullptr ? cap_pre * p.sm_scale : cap_pre;
    float o[NT][MT][4];
    float m[NT][2], l[NT][2];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
#pragma unroll
      for (int i = 0; i < MT; ++i) o[nt][i][0] = o[nt][i][1] = o[nt][i][2] = o[nt][i][3] = 0.f;
      m[nt][0] = m[nt][1] = kNegInf;
      l[nt][0] = l[nt][1] = 0.f;
    }

    for (int tile = sg.tile_begin; tile < sg.tile_end; ++tile) {
      mbar_wait(full_bar + 8 * stage, phase);
      const int tok0 = tile * G::TT + row0;
      const int vcount = min(TPW, sg.tok_hi - tok0);
      if (vcount > 0 && tok0 + TPW > sg.tok_lo) {  // warp-uniform
        const uint32_t kbase = ring + stage * G::STAGE_BYTES;
        const uint32_t vbase = kbase + vdelta;
        if constexpr (LOADER == 0) {
          if (vcount < TPW) {  // rows the TMA did not (validly) fill: zero V so 0 * junk != NaN
            for (int i = lane; i < (TPW - vcount) * G::CHUNKS; i += 32) {
              const int row = row0 + vcount + i / G::CHUNKS;
              const uint32_t a = vbase + StageRow<G::TT>(row, rmask).at(i % G::CHUNKS);
              asm volatile("st.shared.v4.b32 [%0], {%1,%1,%1,%1};" ::"r"(a), "r"(0u) : "memory");
            }
            __syncwarp();
          }
        }
        // S^T (16 tokens x 16 columns): one K fragment per k-step, two MMAs
        float s[NT][4];
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
        {
          const int kch = lane >> 4;
          const StageRow<64> kr(row0 + (lane & 7) + 8 * ((lane >> 3) & 1), rmask);
#pragma unroll
          for (int ks = 0; ks < KS; ++ks) {
            uint32_t a[4];
            ldmatrix_x4(a, kbase + kr.at(2 * ks + kch));
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) mma_16816<BF>(s[nt], a, qf[nt][ks][0], qf[nt][ks][1]);
          }
        }
        const int ta = tok0 + g, tb = ta + 8;
        const bool va = ta >= sg.tok_lo, vb = tb >= sg.tok_lo;
        const float pf[2] = {(float)ta, (float)tb};
        uint32_t pb[NT][2];
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
          bool vc[4];
          vc[0] = va && ta < sg.tok_hi - hid[nt][0]; vc[1] = va && ta < sg.tok_hi - hid[nt][1];
          vc[2] = vb && tb < sg.tok_hi - hid[nt][0]; vc[3] = vb && tb < sg.tok_hi - hid[nt][1];
          float x[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float v = use_cap ? cap_post * tanhf_approx(s[nt][i] * pre + alc[nt][i & 1] * pf[i >> 1])
                                    : s[nt][i] * mul + alc[nt][i & 1] * pf[i >> 1];
            x[i] = vc[i] ? v : kNegInf;
          }
          float mx[2] = {fmaxf(x[0], x[2]), fmaxf(x[1], x[3])};
          float sc[2];
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            mx[j] = fmaxf(mx[j], __shfl_xor_sync(0xffffffffu, mx[j], 4));
            mx[j] = fmaxf(mx[j], __shfl_xor_sync(0xffffffffu, mx[j], 8));
            mx[j] = fmaxf(mx[j], __shfl_xor_sync(0xffffffffu, mx[j], 16));
            const float mn = fmaxf(m[nt][j], mx[j]);
            sc[j] = ex2f(m[nt][j] - mn);
            m[nt][j] = mn;
          }
          float pr[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) pr[i] = vc[i] ? ex2f(x[i] - m[nt][i & 1]) : 0.f;
          l[nt][0] = l[nt][0] * sc[0] + pr[0] + pr[2];
          l[nt][1] = l[nt][1] * sc[1] + pr[1] + pr[3];
#pragma unroll
          for (int i = 0; i < MT; ++i) {
            o[nt][i][0] *= sc[0]; o[nt][i][1] *= sc[1]; o[nt][i][2] *= sc[0]; o[nt][i][3] *= sc[1];
          }
          pb[nt][0] = movmatrix_trans(pack2<BF>(pr[0], pr[1]));
          pb[nt][1] = movmatrix_trans(pack2<BF>(pr[2], pr[3]));
        }
        {  // O^T += V^T . P^T: one V fragment per 16 output dims, two MMAs
          const int vch = (lane >> 3) & 1;
          const StageRow<64> vr(row0 + (lane & 7) + 8 * (lane >> 4), rmask);
#pragma unroll
          for (int mt = 0; mt < MT; ++mt) {
            uint32_t a[4];
            ldmatrix_x4_trans(a, vbase + vr.at(2 * mt + vch));
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) mma_16816<BF>(o[nt][mt], a, pb[nt][0], pb[nt][1]);
          }
        }
        if constexpr (LOADER == 0) {
          if (vcount < TPW) fence_proxy_async_smem();  // generic zero-fill before next TMA write
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(empty_bar + 8 * stage);
      if (++stage == p.num_stages) { stage = 0; phase ^= 1; }
    }

    // ---- segment epilogue: 4 warp states -> one (o, lse) per column; 16 columns in scratch_ml ----
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        l[nt][j] += __shfl_xor_sync(0xffffffffu, l[nt][j], 4);
        l[nt][j] += __shfl_xor_sync(0xffffffffu, l[nt][j], 8);
        l[nt][j] += __shfl_xor_sync(0xffffffffu, l[nt][j], 16);
      }
    if (g == 0) {
#pragma unroll
      for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          scratch_ml[(warp * 16 + 8 * nt + 2 * t + j) * 2 + 0] = m[nt][j];
          scratch_ml[(warp * 16 + 8 * nt + 2 * t + j) * 2 + 1] = l[nt][j];
        }
    }
    if (threadIdx.x == 0) merge_post(p, pm, scratch_ml);  // arrival of the previous partial
    named_bar_sync(1, kConsumerWarps * 32);
    if (threadIdx.x == 0) src.ctl[kQDone0] = k + 1;
    merge_run_posted<QT, HD, true>(p, scratch_ml, threadIdx.x);
    float M[NT][2], L[NT][2];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int c = 8 * nt + 2 * t + j;
        float mm = kNegInf;
#pragma unroll
        for (int w = 0; w < kConsumerWarps; ++w) mm = fmaxf(mm, scratch_ml[(w * 16 + c) * 2]);
        float ll = 0.f;
#pragma unroll
        for (int w = 0; w < kConsumerWarps; ++w)
          ll += scratch_ml[(w * 16 + c) * 2 + 1] * ex2f(scratch_ml[(w * 16 + c) * 2] - mm);
        M[nt][j] = mm; L[nt][j] = ll;
        const float f = ex2f(m[nt][j] - mm);
#pragma unroll
        for (int i = 0; i < MT; ++i) { o[nt][i][j] *= f; o[nt][i][2 + j] *= f; }
      }
    const bool direct = sg.slot < 0;
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {  // the cross-warp sum of o goes through scratch_o one N tile at a time
      if (warp != 0) {
        float* dst = scratch_o + (warp - 1) * (MT * 4 * 32) + lane;
#pragma unroll
        for (int i = 0; i < MT; ++i)
#pragma unroll
          for (int r = 0; r < 4; ++r) dst[(i * 4 + r) * 32] = o[nt][i][r];
      }
      named_bar_sync(2, kConsumerWarps * 32);
      if (warp == 0) {
#pragma unroll
        for (int w = 0; w < kConsumerWarps - 1; ++w) {
          const float* s2 = scratch_o + w * (MT * 4 * 32) + lane;
#pragma unroll
          for (int i = 0; i < MT; ++i)
#pragma unroll
            for (int r = 0; r < 4; ++r) o[nt][i][r] += s2[(i * 4 + r) * 32];
        }
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int c = 8 * nt + 2 * t + j;
          if (c < GS) {
            const RowHead rh = row_head(p, sg.kv_head, c);
            float Lc = L[nt][j];
            const bool any = Lc > 0.f;
            if (direct && any && p.sinks != nullptr)
              Lc += ex2f(__ldg(p.sinks + rh.head) * 1.4426950408889634f - M[nt][j]);
            const float inv = any ? (direct ? p.v_scale : 1.f) / Lc : 0.f;
            const float lse2 = any ? M[nt][j] + lg2f(Lc) : kNegInf;
            if (direct) {
              const int64_t oh = rh.token(p, sg.request) * hq_real + rh.head;
              if (QT == 1) {
                __nv_bfloat16* orow = static_cast<__nv_bfloat16*>(p.o) + oh * HD;
#pragma unroll
                for (int i = 0; i < MT; ++i) {
                  orow[o_dim<0>(i, 0, g)] = __float2bfloat16_rn(o[nt][i][j] * inv);
                  orow[o_dim<0>(i, 1, g)] = __float2bfloat16_rn(o[nt][i][2 + j] * inv);
                }
              } else {
                __half* orow = static_cast<__half*>(p.o) + oh * HD;
#pragma unroll
                for (int i = 0; i < MT; ++i) {
                  orow[o_dim<0>(i, 0, g)] = __float2half_rn(o[nt][i][j] * inv);
                  orow[o_dim<0>(i, 1, g)] = __float2half_rn(o[nt][i][2 + j] * inv);
                }
              }
              if (p.lse != nullptr && g == 0) p.lse[oh] = lse2;
            } else {
              float* orow = p.tmp_o + ((int64_t)sg.slot * GS + c) * HD;
#pragma unroll
              for (int i = 0; i < MT; ++i) {
                orow[o_dim<0>(i, 0, g)] = o[nt][i][j] * inv;
                orow[o_dim<0>(i, 1, g)] = o[nt][i][2 + j] * inv;
              }
              if (g == 0) p.tmp_lse[(int64_t)sg.slot * GS + c] = lse2;
            }
          }
        }
      }
      named_bar_sync(2, kConsumerWarps * 32);  // scratch_o is free for the next N tile / segment
    }
    if (warp == 0 && !direct && sg.merge_unit >= 0) {  // publish the partial: count the arrival, look later
      __syncwarp();
      if (lane == 0) merge_arrive(p, pm, sg);
    }
  }
  if (threadIdx.x == 0) merge_post(p, pm, scratch_ml);
  named_bar_sync(1, kConsumerWarps * 32);
  merge_run_posted<QT, HD, true>(p, scratch_ml, threadIdx.x);
}

// threads per CTA: 4 consumer warps + 1 (TMA) or 4 (cp.async) producer warps
template <int LOADER>
struct BlockDim {
  static constexpr int kProducerWarps = (LOADER == 0) ? 1 : kCpAsyncProducerThreads / 32;
  static constexpr int kThreads = (kConsumerWarps + kProducerWarps) * 32;
};

// QT: 0 f16, 1 bf16 (q / o dtype).  KVF: 0 = KV has q's dtype, 1 = e4m3, 2 = e5m2.
// LOADER: 0 = TMA boxes, 1 = cp.async gathers.
// MROWS: multi-row units (see row_head): n query rows of a request share its KV stream.
// (MROWS as int: 1 = one N tile of 8 columns per unit, 2 = two: consume_rows2.)
template <int QT, int KVF, int HD, int LOADER, bool ROPE = false, int MROWS = 0>
__global__ void __launch_bounds__(BlockDim<LOADER>::kThreads, HD <= 128 ? 2 : 1)
    paged_decode_kernel(const __grid_constant__ DecodeParams p,
                        const __grid_constant__ CUtensorMap kmap,
                        const __grid_constant__ CUtensorMap vmap) {
  using G = TileGeom<KVF, HD>;
  using L = SmemLayout<KVF, HD>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t pad = ((raw + 1023u) & ~1023u) - raw;
  uint8_t* base = smem_raw + pad;
  const uint32_t ring = raw + pad;
  const int ring_bytes = p.num_stages * G::STAGE_BYTES;
  float* scratch_o = reinterpret_cast<float*>(base + ring_bytes);
  float* scratch_ml = reinterpret_cast<float*>(base + ring_bytes + L::SCRATCH_O);
  Segment* seg_cache =
      reinterpret_cast<Segment*>(base + ring_bytes + L::SCRATCH_O + L::SCRATCH_ML);
  const uint32_t full_bar = ring + ring_bytes + L::SCRATCH_O + L::SCRATCH_ML + L::SEG_BYTES;
  const uint32_t empty_bar = full_bar + 8 * kMaxStages;

  const int tid = threadIdx.x;
  FI_TRACE_IF(tid == 0, 0);
#ifdef FI_B200_TRACE
  if (tid == 0) fi_trace_smid();
#endif
  if (tid == 0) {
    for (int s = 0; s < p.num_stages; ++s) {
      mbar_init(full_bar + 8 * s, LOADER == 0 ? 1 : BlockDim<LOADER>::kProducerWarps * 32);
      mbar_init(empty_bar + 8 * s, kConsumerWarps);
    }
    fence_mbar_init();
    if (LOADER == 0) {
      tma_prefetch_desc(&kmap);
      tma_prefetch_desc(&vmap);
    }
  }
  // the work list was written by plan() (a memcpy), not by the preceding kernel, so it
  // may be read before the PDL dependency resolves
  const SegSource src = seg_source_init(p, G::TT, seg_cache, scratch_ml, LOADER != 0, LOADER != 0);
  __syncthreads();
  FI_TRACE_IF(tid == 0, 1);
  if (seg_source_idle(src)) return;
  // Early trigger: the next kernel in the stream becomes launchable now, so its CTAs move into
  // SM slots as ours retire (they run their prologue and then block in their own
  // griddepcontrol.wait until this grid has completed and flushed) — the launch latency and
  // the tail imbalance of this grid overlap the ramp of the next one.
  if (p.enable_pdl == 1) pdl_launch_dependents();
  FI_TRACE_IF(tid == 0, 2);

  // every role waits for the upstream kernel itself, right before its first access to data an
  // upstream kernel may have produced (q, the KV cache, counters)
  // cp.async loader with two CTAs per SM: the kernel is compiled for 128 registers per thread
  // (256 threads x 2 CTAs); the consumer warpgroup grows to 184, the producer warpgroup shrinks to 72
  constexpr bool kRegSplit = LOADER != 0 && HD <= 128;
  if (tid < kConsumerWarps * 32) {
    if constexpr (kRegSplit) setmaxnreg_inc<kCpAsyncConsumerRegs>();
    if (p.enable_pdl) pdl_wait();
    if constexpr (MROWS == 2) consume_rows2<QT, HD, LOADER>(p, src, ring, full_bar, empty_bar, scratch_o, scratch_ml);
    else consume<QT, KVF, HD, LOADER, ROPE, MROWS == 1>(p, src, ring, full_bar, empty_bar, scratch_o, scratch_ml);
  } else if (LOADER == 0) {
    produce_tma<KVF, HD, true>(p, src, &kmap, &vmap, ring, full_bar, empty_bar);  // waits inside
  } else {
    if constexpr (kRegSplit) setmaxnreg_dec<kCpAsyncProducerRegs>();
    if (p.enable_pdl) pdl_wait();
    produce_cpasync<KVF, HD>(p, src, tid - kConsumerWarps * 32, ring, full_bar, empty_bar);
  }
  if (p.enable_pdl == 2) pdl_launch_dependents();  // FI_B200_PDL_LATE=1 (A/B runs)
}

}  // namespace fib200
