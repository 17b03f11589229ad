// plan.cu — host-side decode scheduler (C++), the replacement for the reference's
// DecodePlan (include/flashinfer/attention/scheduler.cuh:424-493) and its helpers
// PartitionPagedKVCacheBinarySearchMinNumPagePerBatch (:73-99) and DecodeSplitKVIndptr
// (:348-364).
//
// The reference picks ONE uniform chunk size for all requests and launches a
// (work item, kv_head) grid.  Here the KV tiles (64 tokens) of all (request, kv_head)
// units are laid out as one stream, request-major / head-minor, and the stream is cut
// into `num_ctas` equal contiguous ranges — one per persistent CTA — so every CTA reads
// the same number of KV bytes (+-1 tile) whatever the length distribution.  A unit cut
// by a range boundary becomes several partial segments (consecutive partial slots) plus
// one MergeUnit; a unit inside one range is written directly.
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <mutex>
#include <vector>

#include "host_common.h"

namespace fib200 {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

static int env_int(const char* name, int dflt, int lo, int hi);

// process-wide overrides: fi_set_tuning(), initialised from FI_B200_STAGES / FI_B200_CTAS_PER_SM
// (A/B runs of bench.py without touching its code)
Tuning& tuning() {
  static Tuning t = {env_int("FI_B200_STAGES", 0, 0, kMaxStages), env_int("FI_B200_CTAS_PER_SM", 0, 0, 2), 0, 0};
  return t;
}

KernelConfig choose_config(int kvf, int head_dim) {
  const int es = kvf == 0 ? 2 : 1;
  const int stage = 2 * tile_tokens(kvf) * head_dim * es;
  const int fixed = 1024 + (kConsumerWarps - 1) * head_dim * 32 /*scratch o*/ +
                    kScratchMlBytes + kSegCache * (int)sizeof(Segment) + 2 * kMaxStages * 8 + 128;
  const int sm_total = 233472;  // 228 KiB per SM, each resident CTA reserves 1 KiB
  KernelConfig c;
  c.ctas_per_sm = tuning().ctas_per_sm;
  if (c.ctas_per_sm <= 0) {
    const int avail2 = sm_total / 2 - 1024 - fixed;
    c.ctas_per_sm = (avail2 / stage >= 3) ? 2 : 1;
  }
  const int avail = sm_total / c.ctas_per_sm - 1024 - fixed;
  int stages = std::min(kMaxStages, avail / stage);
  if (tuning().stages > 0) stages = std::min(stages, tuning().stages);
  c.stages = std::max(stages, 1);
  c.smem_bytes = fixed + c.stages * stage;
  return c;
}

int device_sm_count(int device, int* out) {
  static std::mutex mu;
  static int cache[64];
  std::lock_guard<std::mutex> lk(mu);
  if (device < 0 || device >= 64) return fail(FI_ERR_INVALID_ARGUMENT, "bad device %d", device);
  if (cache[device] == 0) {
    int n = 0;
    FI_CUDA_TRY(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device));
    cache[device] = n;
  }
  *out = cache[device];
  return FI_OK;
}

static inline size_t align16(size_t x) { return (x + 15) & ~size_t(15); }

static int env_int(const char* name, int dflt, int lo, int hi) {
  const char* e = getenv(name);
  if (e == nullptr || e[0] == 0) return dflt;
  const int v = atoi(e);
  return std::min(hi, std::max(lo, v));
}

// Head-grouped stream order, on by default; FI_B200_HEAD_GROUP=0 restores the request-major /
// head-minor order (read at every plan so that A/B runs fit in one process).  Measured on B200:
// C2 170.0 -> 164.4 us, C3 (e4m3, 128-byte rows) 397 -> 376 us, C4 1401 -> 1349 us.
static bool head_group_enabled() {
  const char* e = getenv("FI_B200_HEAD_GROUP");
  return e != nullptr ? e[0] != '0' : true;
}

const SchedTuning& sched_tuning() {
  static const SchedTuning t = [] {
    SchedTuning s;
    // default: the static equal split.  Guided self-scheduling (e.g. F0=90 G=50 TAIL=4) was
    // measured on B200 within +-2 % of it on 1 GB shapes and slower on small ones — every extra
    // segment boundary costs more than the tail imbalance it removes.
    s.f0_pct = env_int("FI_B200_SCHED_F0", 100, 1, 100);
    s.g_pct = env_int("FI_B200_SCHED_G", 50, 5, 100);
    s.tail_tiles = env_int("FI_B200_SCHED_TAIL", 4, 1, 1 << 20);
    s.lookahead_tiles = env_int("FI_B200_SCHED_LOOKAHEAD", 4, 0, 1 << 20);
    return s;
  }();
  return t;
}

}  // namespace fib200

using namespace fib200;

extern "C" {

const char* fi_last_error(void) { return g_err; }
int fi_version(void) { return 100; }
int fi_plan_info_len(void) { return FI_PLAN_INFO_LEN; }

void fi_set_tuning(int stages, int ctas_per_sm, int loader, int min_tiles_per_cta) {
  Tuning& t = tuning();
  t.stages = stages;
  t.ctas_per_sm = ctas_per_sm;
  t.loader = loader;
  t.min_tiles_per_cta = min_tiles_per_cta;
}

// FI_B200_HEAD_PAIRS=0 ignores FI_PLAN_HEAD_PAIRS (A/B runs; read at every plan)
static bool head_pairs_enabled() {
  const char* e = getenv("FI_B200_HEAD_PAIRS");
  if (e != nullptr && e[0] == '0') return false;
  e = getenv("FI_B200_FP8_TC5");
  return !(e != nullptr && e[0] == '0');
}

static int plan_impl(void* float_ws, size_t float_ws_bytes, void* int_ws, void* pinned_int_ws,
                     size_t int_ws_bytes, const int32_t* indptr_host,
                     const int32_t* last_page_len_host, const int32_t* kv_len_host,
                     int batch_size, int num_qo_heads, int num_kv_heads, int head_dim,
                     int page_size, int q_dtype, int kv_dtype, int window_left,
                     int cuda_graph, int fixed_split_pages, int disable_split_kv, int device,
                     int num_ctas_override, void* stream, const int32_t* kv_indices_host,
                     int64_t num_indices, int64_t* plan_info_out, int flags) {
  static_assert(kPiLen <= FI_PLAN_INFO_LEN, "plan info vector too short");
  (void)float_ws;
  if (batch_size < 0 || num_qo_heads <= 0 || num_kv_heads <= 0 || page_size <= 0)
    return fail(FI_ERR_INVALID_ARGUMENT, "invalid problem shape");
  if (num_qo_heads % num_kv_heads != 0)
    return fail(FI_ERR_INVALID_ARGUMENT, "num_qo_heads %d is not a multiple of num_kv_heads %d",
                num_qo_heads, num_kv_heads);
  // a GQA group larger than one MMA N tile becomes kv_div virtual KV heads per real one (the
  // reference's CUDA-core decode stops at group 8; its tensor-core decode takes any group)
  const int kvf0 = kv_format(q_dtype, kv_dtype);
  // FI_PLAN_WIDE_UNITS: 16 columns per unit (two MMA N tiles, decode_kernel.cuh consume_rows2)
  const char* wide_env = getenv("FI_B200_WIDE_UNITS");  // 0: ignore the flag (A/B runs; read at every plan)
  const char* tc5_env = getenv("FI_B200_FP8_TC5");
  const bool wide_shape = kvf0 == 0 ? head_dim <= 128
                                    : (kvf0 > 0 && head_dim == 128 && !(tc5_env != nullptr && tc5_env[0] == '0'));
  const bool wide_units = (flags & FI_PLAN_WIDE_UNITS) != 0 && wide_shape && page_size >= 8 &&
                          (page_size & (page_size - 1)) == 0 && tuning().loader != 2 &&
                          !(wide_env != nullptr && wide_env[0] == '0');
  const int kv_div = group_split(num_qo_heads / num_kv_heads, wide_units ? 2 * kMaxGroup : max_group_for(head_dim));
  num_kv_heads *= kv_div;
  int group = num_qo_heads / num_kv_heads;
  const int kvf = kv_format(q_dtype, kv_dtype);
  if (kvf < 0) return fail(FI_ERR_UNSUPPORTED, "unsupported q/kv dtype pair (%d, %d)", q_dtype, kv_dtype);
  // Head-pair plan (FI_PLAN_HEAD_PAIRS, fp8 KV, head_dim 128, TMA-sized pages): a unit is a request and a
  // PAIR of adjacent KV heads, a tile 64 tokens of both — in an NHD cache the two 128-byte rows of a
  // token are adjacent, so the tcgen05 kernel streams 256 contiguous bytes per token like a 16-bit
  // cache does.  To everything downstream (items, partial slots, merge) this is a problem with half
  // the KV heads and twice the GQA group.
  const bool head_pairs = (flags & FI_PLAN_HEAD_PAIRS) != 0 && head_pairs_enabled() && kvf != 0 &&
                          head_dim == 128 && kv_div == 1 && num_kv_heads % 2 == 0 && group <= kMaxGroup &&
                          page_size >= 8 && (page_size & (page_size - 1)) == 0 && tuning().loader != 2;
  if (head_pairs) {
    num_kv_heads /= 2;
    group *= 2;
  }
  // fp8 KV at head_dim 64 (64-byte rows, below the 128-byte swizzle atom of the MMA kernels) and head_dim
  // 512 run on the CUDA-core variant kernel (decode_rope_kernel.cuh with zero rotation frequencies)
  if (!(head_dim == 64 || head_dim == 128 || head_dim == 256 || head_dim == 512))
    return fail(FI_ERR_UNSUPPORTED, "unsupported head_dim %d for this dtype", head_dim);
  if (indptr_host == nullptr || (last_page_len_host == nullptr && kv_len_host == nullptr))
    return fail(FI_ERR_INVALID_ARGUMENT, "indptr / last_page_len must be host pointers");
  if (pinned_int_ws == nullptr || int_ws == nullptr)
    return fail(FI_ERR_INVALID_ARGUMENT, "int workspaces must not be null");

  const KernelConfig cfg = choose_config(kvf, head_dim);
  int num_ctas = num_ctas_override;
  if (num_ctas <= 0) {
    int sms = 0;
    int rc = device_sm_count(device, &sms);
    if (rc != FI_OK) return rc;
    num_ctas = sms * cfg.ctas_per_sm;
  }

  // ---- per-request token / tile ranges --------------------------------------
  const int TT = head_pairs ? tile_tokens(kvf) / 2 : tile_tokens(kvf);
  // one scratch allocation per thread, reused across plans (plan() runs once per decode step)
  static thread_local std::vector<int32_t> scratch;
  scratch.resize((size_t)4 * batch_size);
  int32_t* tok_lo = scratch.data();
  int32_t* tok_hi = tok_lo + batch_size;
  int32_t* tile_b = tok_hi + batch_size;
  int32_t* ntile = tile_b + batch_size;
  int64_t total_tiles = 0;
  for (int b = 0; b < batch_size; ++b) {
    const int64_t npages = (int64_t)indptr_host[b + 1] - indptr_host[b];
    if (npages < 0) return fail(FI_ERR_INVALID_ARGUMENT, "indptr must be non-decreasing");
    int64_t len;
    if (kv_len_host != nullptr) {
      len = kv_len_host[b];
      if (len < 0 || len > npages * page_size)
        return fail(FI_ERR_INVALID_ARGUMENT, "kv_len[%d]=%lld exceeds its pages", b, (long long)len);
    } else {
      const int lpl = last_page_len_host[b];
      if (npages > 0 && (lpl < 0 || lpl > page_size))
        return fail(FI_ERR_INVALID_ARGUMENT, "last_page_len[%d]=%d out of range", b, lpl);
      len = npages == 0 ? 0 : (npages - 1) * page_size + lpl;
    }
    if (len > 0x7fffff00LL) return fail(FI_ERR_INVALID_ARGUMENT, "kv_len too large");
    int64_t lo = 0;
    if (window_left >= 0) lo = std::max<int64_t>(0, len - 1 - window_left);
    tok_lo[b] = (int32_t)lo;
    tok_hi[b] = (int32_t)len;
    tile_b[b] = (int32_t)(lo / TT);
    ntile[b] = len == 0 ? 0 : (int32_t)((len + TT - 1) / TT - tile_b[b]);
  }
  // Batch-invariant modes (reference decode.py:880-887): with fixed_split_size every request is
  // cut at multiples of S tokens from its start, each chunk is ONE indivisible segment whose
  // arithmetic does not depend on what else is in the batch, and a request with more than one
  // chunk is always merged from its chunk partials in chunk order; with disable_split_kv a unit
  // is never cut at all.  Item boundaries then fall only between chunks (coarser balance).
  const int64_t S = fixed_split_pages > 0 ? (int64_t)fixed_split_pages * page_size : 0;
  const bool indivisible = S > 0 || disable_split_kv != 0;
  auto chunk_count = [&](int b) -> int64_t {
    if (ntile[b] == 0) return 0;
    if (S <= 0) return 1;
    return ((int64_t)tok_hi[b] + S - 1) / S - tok_lo[b] / S;
  };
  auto chunk_bounds = [&](int b, int64_t c, int32_t& lo, int32_t& hi) {
    if (S <= 0) { lo = tok_lo[b]; hi = tok_hi[b]; return; }
    const int64_t c0 = tok_lo[b] / S + c;
    lo = (int32_t)std::max<int64_t>(tok_lo[b], c0 * S);
    hi = (int32_t)std::min<int64_t>(tok_hi[b], (c0 + 1) * S);
  };
  for (int b = 0; b < batch_size; ++b) {
    const int64_t nc = chunk_count(b);
    int64_t t = 0;
    for (int64_t c = 0; c < nc; ++c) {  // a tile shared by two chunks is streamed by both
      int32_t lo, hi;
      chunk_bounds(b, c, lo, hi);
      t += (hi + TT - 1) / TT - lo / TT;
    }
    total_tiles += t * num_kv_heads;
  }
  if (total_tiles > 0x7fffffffLL) return fail(FI_ERR_INVALID_ARGUMENT, "too many KV tiles");

  // ---- work items: guided self-scheduling over the tile stream ------------------------
  // Round 0 hands every CTA one item of f0 of its fair share (the kernel starts CTA i on
  // item i without a ticket); every later round cuts g of what is left into num_ctas equal
  // items, down to `tail` tiles per item.  CTAs draw those in stream order, so an SM that
  // streams faster (HBM / L2 arbitration is not uniform across the two dies) ends up with
  // more of the tail instead of idling.  f0 = 100 reproduces the static equal split.
  const SchedTuning st = sched_tuning();
  // smallest item: 2 tiles (a bs = 1, kv 2048 decode takes 8.7 us with 2-tile items, 10.2 us with 4)
  const int min_tiles = tuning().min_tiles_per_cta > 0 ? tuning().min_tiles_per_cta : 2;
  // Uniform batches (every non-empty request has the same tile count n_u): a chunk that divides n_u
  // — or is a multiple of it — keeps every CTA inside one unit, i.e. one segment epilogue per CTA
  // and no unit cut in odd places; taken when it costs at most 1/6 more tiles per CTA
  // (bs 4 x kv 4096: 296 items of 7 tiles 18.6 us -> 256 items of 8 tiles 15.9 us; 8-way per-rank
  // C2: 293 items of 14 tiles 27.3 us -> 256 items of 16 tiles 25.6 us; per-rank C3 53.0 -> 50.8 us).
  int64_t uniform_tiles = -1;
  for (int b = 0; b < batch_size; ++b) {
    if (ntile[b] == 0) continue;
    if (uniform_tiles < 0) uniform_tiles = ntile[b];
    else if (uniform_tiles != ntile[b]) { uniform_tiles = 0; break; }
  }
  auto align_chunk = [&](int64_t c0) -> int64_t {
    const int64_t n_u = uniform_tiles;
    if (n_u <= 0 || indivisible || st.f0_pct < 100 || c0 <= 0) return c0;
    static const int den_short = env_int("FI_B200_ALIGN_DEN", 6, 1, 1 << 20);  // A/B: 1000 = never align
    static const int den_long = env_int("FI_B200_ALIGN_DEN_LONG", 32, 1, 1 << 20);
    // Aligned items leave CTA slots idle (C2 / C3: 256 items on 296 slots; head_dim 256, one CTA per SM:
    // 128 items on 148 SMs).  That pays for short items, where a cut unit costs a merge chain as long as
    // the item itself, and for 16-bit KV at head_dim 128, which saturates DRAM with 256 CTAs (C2 158.6 vs
    // 159.3 us).  Every other kernel turns idle slots into lost bytes in flight: with items of >= 2 MiB
    // they round by <= 1/32 only — C3 on one B200 351.7 -> 326.2 us (NHD, head pairs), 343.6 -> 325.8 us
    // (HND); 1 GiB probes: bf16 D = 64 209.4 -> 185.6 us, D = 256 166.6 -> 162.5 us, e4m3 D = 256
    // 264.9 -> 240.9 us.
    const int64_t tile_bytes = 2LL * TT * head_dim * (kvf == 0 ? 2 : 1) * (head_pairs ? 2 : 1);
    const bool long_items = c0 * tile_bytes >= (2LL << 20) && !(kvf == 0 && head_dim == 128);
    const int den = long_items ? den_long : den_short;
    const int64_t hi = c0 + std::max<int64_t>(long_items ? 0 : 1, c0 / den);
    if (c0 >= n_u) {
      const int64_t c = (c0 + n_u - 1) / n_u * n_u;
      return c <= hi ? c : c0;
    }
    for (int64_t c = c0; c <= hi; ++c)
      if (n_u % c == 0) return c;
    return c0;
  };
  // Head-grouped order (static divisible plans, Hkv > 1): the REQUEST tile stream is cut into
  // num_ctas / Hkv ranges and the Hkv CTAs of a group walk the same range, one KV head each, in
  // step.  In an NHD cache the Hkv head rows of a token are adjacent in memory, so the group
  // sweeps whole pages at about the same time (DRAM row locality) instead of a 1-in-Hkv comb.
  const bool head_group = head_group_enabled() && !indivisible && num_kv_heads > 1 &&
                          num_ctas % num_kv_heads == 0 && st.f0_pct >= 100 && total_tiles > 0;
  static thread_local std::vector<int64_t> item_end;  // exclusive end (stream position) of every item
  item_end.clear();
  if (!head_group) {  // (the head-grouped walk below has its own, regular item boundaries)
    int64_t pos = 0;
    bool first = true;
    while (pos < total_tiles) {
      const int64_t rem = total_tiles - pos;
      int64_t chunk = first ? (total_tiles * st.f0_pct + 100LL * num_ctas - 1) / (100LL * num_ctas)
                            : (rem * st.g_pct + 100LL * num_ctas - 1) / (100LL * num_ctas);
      if (first && (st.f0_pct >= 100 || indivisible)) chunk = (total_tiles + num_ctas - 1) / num_ctas;
      chunk = std::max<int64_t>(chunk, first ? min_tiles : st.tail_tiles);
      if (first) chunk = align_chunk(chunk);
      for (int i = 0; i < num_ctas && pos < total_tiles; ++i) {
        pos = std::min(total_tiles, pos + chunk);
        item_end.push_back(pos);
      }
      first = false;
    }
    if (item_end.empty()) item_end.push_back(0);  // only zero-tile segments (or none at all)
  }
  int64_t hg_chunk = 0, hg_groups = 0;
  if (head_group) {
    const int64_t tr = total_tiles / num_kv_heads, groups = num_ctas / num_kv_heads;
    hg_chunk = align_chunk(std::max<int64_t>((tr + groups - 1) / groups, min_tiles));
    hg_groups = (tr + hg_chunk - 1) / hg_chunk;
  }
  const int64_t num_items = head_group ? hg_groups * num_kv_heads : (int64_t)item_end.size();
  const int64_t* const item_end_p = item_end.data();  // not through the TLS wrapper in the hot loop

  // ---- layout: offsets of everything the kernels get as launch parameters depend only on
  // (bs, heads) — CUDA-graph replays of an earlier run() stay valid after a re-plan; the two
  // variable-length arrays (item heads, segment list) come last -----------------------------
  const int64_t units = (int64_t)batch_size * num_kv_heads;
  // every item boundary adds at most one segment; fixed-split chunks are one segment each
  int64_t max_segs = units + num_items;
  if (S > 0) {
    max_segs = 0;
    for (int b = 0; b < batch_size; ++b) max_segs += std::max<int64_t>(chunk_count(b), 1) * num_kv_heads;
  }
  const int64_t max_merge = std::max<int64_t>(units, 1);
  size_t off = 0;
  const size_t off_header = off;   off = align16(off + sizeof(PlanHeader));
  const size_t off_work = off;     off = align16(off + 16);
  const size_t off_counters = off; off = align16(off + (size_t)max_merge * 4);
  const size_t off_merge = off;    off = align16(off + (size_t)max_merge * sizeof(MergeUnit));
  // every launched CTA reads head blockIdx.x: heads past the last item are empty
  const int64_t head_cap = std::max<int64_t>(num_items, num_ctas);
  const size_t off_heads = off;    off = align16(off + (size_t)head_cap * sizeof(ItemHead));
  const size_t off_segs = off;     off = align16(off + (size_t)max_segs * sizeof(Segment));
  const size_t int_need = off;
  // The page table, when plan() has it on the host: appended to the work list so that ONE H2D copy
  // brings everything run() needs (outside CUDA-graph mode, where the caller's fixed buffer is the
  // kernel parameter), and the page ids of every item's first tiles go into its head.
  if (kv_indices_host != nullptr) {
    if (num_indices < 0 || (batch_size > 0 && indptr_host[batch_size] > num_indices))
      return fail(FI_ERR_INVALID_ARGUMENT, "kv_indices has %lld entries, indptr needs %d",
                  (long long)num_indices, batch_size > 0 ? indptr_host[batch_size] : 0);
  }
  const bool idx_in_ws = kv_indices_host != nullptr && !cuda_graph &&
                         off + (size_t)num_indices * 4 + 16 <= int_ws_bytes;
  const size_t off_idx = off;
  if (int_need > int_ws_bytes)
    return fail(FI_ERR_WORKSPACE,
                "Buffer overflow when allocating the decode work list: need %zu bytes of int "
                "workspace, have %zu. Increase the workspace buffer size.", int_need, int_ws_bytes);
  // float workspace: [lse of every slot | o of every slot]; the split point depends only on
  // the workspace size and (group, head_dim), never on the plan
  const size_t slot_cap = float_ws_bytes / ((size_t)group * (head_dim + 1) * 4 + 1);
  const size_t off_tmp_lse = 0;
  const size_t off_tmp_o = align16(slot_cap * group * 4);

  char* host = static_cast<char*>(pinned_int_ws);
  PlanHeader* hdr = reinterpret_cast<PlanHeader*>(host + off_header);
  Segment* segs = reinterpret_cast<Segment*>(host + off_segs);
  MergeUnit* merges = reinterpret_cast<MergeUnit*>(host + off_merge);
  ItemHead* heads = reinterpret_cast<ItemHead*>(host + off_heads);
  memset(host + off_work, 0, 16);

  // split-KV merge: fused into the decode kernel (default) or a second launch
  // (FI_B200_FUSED_MERGE=0, kept for A/B measurements)
  static const bool fused_merge = [] {
    const char* fm = getenv("FI_B200_FUSED_MERGE");
    return !(fm != nullptr && fm[0] == '0');
  }();

  // ---- walk the units in stream order and cut them at the item boundaries -----------------
  int64_t nseg = 0, nmerge = 0, nslot = 0;
  int64_t item = 0;  // item currently being filled
  heads[0].first = 0;
  int64_t pos = 0;   // stream position (tiles)
  // A segment is appended to the list and, when it is one of the first kItemInlineSegs of its
  // item, also copied into the item head while it is still in registers.
  auto push = [&](const Segment& sg) {
    const int64_t k = nseg - heads[item].first;
    if (k < kItemInlineSegs) heads[item].seg[k] = sg;
    segs[nseg++] = sg;
  };
  auto own = [&](int64_t at) {  // advance to the item that owns stream position `at`
    while (item_end_p[item] <= at) {
      ++item;
      heads[item].first = (int32_t)nseg;
    }
  };
  if (head_group) {
    // per request: position in the request tile stream, first / last group, merge + slot bases
    static thread_local std::vector<int64_t> hg;
    hg.resize((size_t)5 * batch_size);
    int64_t* pos_b = hg.data();
    int64_t* g_first = pos_b + batch_size;
    int64_t* g_last = g_first + batch_size;
    int64_t* mbase = g_last + batch_size;
    int64_t* sbase = mbase + batch_size;
    int64_t at = 0;
    for (int b = 0; b < batch_size; ++b) {
      pos_b[b] = at;
      if (ntile[b] == 0) {
        g_first[b] = g_last[b] = std::min(at / hg_chunk, hg_groups - 1);
      } else {
        g_first[b] = at / hg_chunk;
        g_last[b] = (at + ntile[b] - 1) / hg_chunk;
      }
      at += ntile[b];
      const int64_t pieces = g_last[b] - g_first[b] + 1;
      mbase[b] = -1; sbase[b] = 0;
      if (pieces > 1) {
        mbase[b] = nmerge; sbase[b] = nslot;
        for (int h = 0; h < num_kv_heads; ++h) {
          MergeUnit mu = {b, h, (int32_t)(nslot + h * pieces), (int32_t)pieces};
          merges[nmerge++] = mu;
        }
        nslot += pieces * num_kv_heads;
      }
    }
    // Every head of a group walks the same request range: the group's segments are built ONCE
    // (as for head 0) and replicated per head with the head-dependent fields patched — the walk
    // is the planner's hot loop (C2: 512 segments).
    static thread_local std::vector<Segment> tmpl;
    static thread_local std::vector<int32_t> tmpl_pieces;
    int b_lo = 0;
    for (int64_t g = 0; g < hg_groups; ++g) {
      const int64_t gs = g * hg_chunk, ge = gs + hg_chunk;
      while (b_lo < batch_size && g_last[b_lo] < g) ++b_lo;
      tmpl.clear();
      tmpl_pieces.clear();
      for (int b = b_lo; b < batch_size && g_first[b] <= g; ++b) {
        Segment s;
        s.request = b; s.kv_head = 0; s.tok_lo = tok_lo[b]; s.tok_hi = tok_hi[b];
        s.page_base = indptr_host[b]; s.kv_len = tok_hi[b];
        s.slot = -1; s.merge_unit = -1; s.merge_count = 0; s.merge_slot = 0;
        const int64_t lo = std::max(gs, pos_b[b]), hi = std::min(ge, pos_b[b] + ntile[b]);
        s.tile_begin = ntile[b] == 0 ? 0 : tile_b[b] + (int32_t)(lo - pos_b[b]);
        s.tile_end = ntile[b] == 0 ? 0 : tile_b[b] + (int32_t)(hi - pos_b[b]);
        int32_t pieces = 0;
        if (mbase[b] >= 0) {
          pieces = (int32_t)(g_last[b] - g_first[b] + 1);
          s.merge_unit = fused_merge ? (int32_t)mbase[b] : -1;
          s.merge_count = pieces;
          s.merge_slot = (int32_t)sbase[b];
          s.slot = s.merge_slot + (int32_t)(g - g_first[b]);
        }
        tmpl.push_back(s);
        tmpl_pieces.push_back(pieces);
      }
      const int nt = (int)tmpl.size();
      const Segment* tp = tmpl.data();
      const int32_t* pp = tmpl_pieces.data();
      for (int h = 0; h < num_kv_heads; ++h) {
        item = g * num_kv_heads + h;
        ItemHead& hd = heads[item];
        hd.first = (int32_t)nseg;
        Segment* dst = segs + nseg;
        for (int i = 0; i < nt; ++i) {
          Segment s = tp[i];
          s.kv_head = h;
          if (pp[i] > 0) {  // head h of a split request: its own merge unit and slot range
            if (fused_merge) s.merge_unit += h;
            s.merge_slot += h * pp[i];
            s.slot += h * pp[i];
          }
          dst[i] = s;
          if (i < kItemInlineSegs) hd.seg[i] = s;
        }
        nseg += nt;
      }
    }
  }
  for (int b = 0; b < batch_size && !head_group; ++b) {
    const int64_t nc = chunk_count(b);
    for (int h = 0; h < num_kv_heads; ++h) {
      Segment s;
      s.request = b; s.kv_head = h; s.tok_lo = tok_lo[b]; s.tok_hi = tok_hi[b];
      s.page_base = indptr_host[b]; s.kv_len = tok_hi[b];
      s.slot = -1; s.merge_unit = -1; s.merge_count = 0; s.merge_slot = 0;
      if (nc == 0) {  // empty request: zero-tile segment, writes o = 0, lse = -5e4
        s.tile_begin = s.tile_end = 0;
        push(s);
        continue;
      }
      // how many segments will this unit become?  (known up front, so that every partial can
      // carry its unit's count / first slot and nothing has to be patched afterwards)
      int64_t pieces = nc;
      if (!indivisible) {
        own(pos);
        const int64_t end = pos + ntile[b];
        int64_t k = item;
        while (item_end_p[k] < end) ++k;
        pieces = k - item + 1;
      }
      const bool split = pieces > 1;
      if (split) {
        MergeUnit mu = {b, h, (int32_t)nslot, (int32_t)pieces};
        merges[nmerge] = mu;
        s.merge_unit = fused_merge ? (int32_t)nmerge : -1;
        s.merge_count = (int32_t)pieces;
        s.merge_slot = (int32_t)nslot;
        ++nmerge;
      }
      for (int64_t c = 0; c < nc; ++c) {
        int32_t lo, hi;
        chunk_bounds(b, c, lo, hi);
        const int32_t tb = lo / TT, te = (hi + TT - 1) / TT;
        const int64_t end = pos + (te - tb);
        own(pos);
        if (indivisible) {  // the chunk is one segment of the item it starts in
          s.tok_lo = lo; s.tok_hi = hi;
          s.tile_begin = tb; s.tile_end = te;
          if (split) s.slot = (int32_t)nslot++;
          push(s);
        } else {  // cut the unit at the item boundaries
          int64_t cur = pos;
          while (cur < end) {
            own(cur);
            const int64_t e = std::min(end, item_end_p[item]);
            s.tile_begin = tb + (int32_t)(cur - pos);
            s.tile_end = tb + (int32_t)(e - pos);
            if (split) s.slot = (int32_t)nslot++;
            push(s);
            cur = e;
          }
        }
        pos = end;
      }
    }
  }
  for (int64_t i = item + 1; i < num_items; ++i) heads[i].first = (int32_t)nseg;  // never reached
  if (item + 1 > num_items || nseg > max_segs || nmerge > max_merge)
    return fail(FI_ERR_WORKSPACE, "internal planner overflow (%lld items, %lld segs, %lld merges)",
                (long long)item + 1, (long long)nseg, (long long)nmerge);
  if ((size_t)nslot > slot_cap)
    return fail(FI_ERR_WORKSPACE,
                "Buffer overflow when allocating split-KV partials: need %zu bytes of float "
                "workspace, have %zu. Increase the workspace buffer size.",
                (size_t)nslot * group * (head_dim + 1) * 4 + 64, float_ws_bytes);
  // item heads: segment range + the first segments inline
  for (int64_t i = num_items; i < head_cap; ++i) heads[i].first = (int32_t)nseg;
  for (int64_t i = 0; i < head_cap; ++i) {
    ItemHead& hd = heads[i];
    hd.count = (int32_t)((i + 1 < num_items ? heads[i + 1].first : nseg) - hd.first);
    hd.num_items = (int32_t)num_items;
    hd.num_pages = 0;
    if (kv_indices_host != nullptr && hd.count > 0 && hd.seg[0].tile_end > hd.seg[0].tile_begin) {
      const Segment& s0 = hd.seg[0];
      const int64_t first_page = (int64_t)s0.tile_begin * TT / page_size;
      const int64_t req_pages = (int64_t)indptr_host[s0.request + 1] - indptr_host[s0.request];
      const int64_t n = std::min<int64_t>(kItemInlinePages, req_pages - first_page);
      if (n > 0) memcpy(hd.pages, kv_indices_host + s0.page_base + first_page, (size_t)n * sizeof(int32_t));
      hd.num_pages = (int32_t)std::max<int64_t>(n, 0);
    }
  }
  // arrival counters of the fused merge: zeroed by every plan (the workspace is not guaranteed
  // to be zero-initialised) and returned to zero by the kernel itself
  memset(host + off_counters, 0, (size_t)std::max<int64_t>(nmerge, 1) * 4);
  hdr->num_items = (int32_t)num_items;
  hdr->num_segments = (int32_t)nseg;
  hdr->num_merge_units = (int32_t)nmerge;
  hdr->num_slots = (int32_t)nslot;
  hdr->total_tiles = (int32_t)total_tiles;
  hdr->pad[0] = hdr->pad[1] = hdr->pad[2] = 0;

  // ---- H2D: [header, work counter, counters(nmerge)] [merges(nmerge)] [heads, segs(nseg)] —
  // as one copy when the unused capacity in between is small ------------------------------
  const size_t used_segs = off_segs + (size_t)std::max<int64_t>(nseg, 1) * sizeof(Segment);
  size_t used = used_segs;
  if (idx_in_ws) {
    memcpy(host + off_idx, kv_indices_host, (size_t)num_indices * 4);
    used = off_idx + (size_t)num_indices * 4;
  }
  if (device >= 0) {
    FI_CUDA_TRY(cudaSetDevice(device));
    cudaStream_t cs = static_cast<cudaStream_t>(stream);
    char* dev = static_cast<char*>(int_ws);
    const size_t gap = (size_t)(max_merge - nmerge) * (4 + sizeof(MergeUnit)) +
                       (idx_in_ws ? off_idx - used_segs : 0);
    if (gap <= (64u << 10)) {
      FI_CUDA_TRY(cudaMemcpyAsync(dev, host, used, cudaMemcpyHostToDevice, cs));
    } else {
      const size_t n0 = off_counters + (size_t)nmerge * 4;
      FI_CUDA_TRY(cudaMemcpyAsync(dev, host, n0, cudaMemcpyHostToDevice, cs));
      if (nmerge > 0)
        FI_CUDA_TRY(cudaMemcpyAsync(dev + off_merge, host + off_merge, (size_t)nmerge * sizeof(MergeUnit),
                                    cudaMemcpyHostToDevice, cs));
      FI_CUDA_TRY(cudaMemcpyAsync(dev + off_heads, host + off_heads, used_segs - off_heads,
                                  cudaMemcpyHostToDevice, cs));
      if (idx_in_ws && num_indices > 0)
        FI_CUDA_TRY(cudaMemcpyAsync(dev + off_idx, host + off_idx, (size_t)num_indices * 4,
                                    cudaMemcpyHostToDevice, cs));
    }
  } else {
    if (int_ws != pinned_int_ws) memcpy(int_ws, pinned_int_ws, used);  // CPU-only test mode
  }

  int64_t* pi = plan_info_out;
  for (int i = 0; i < FI_PLAN_INFO_LEN; ++i) pi[i] = 0;
  pi[kPiMagic] = kPlanMagic;
  pi[kPiNumCtas] = num_ctas;
  pi[kPiNumSegments] = nseg;
  pi[kPiNumMergeUnits] = nmerge;
  pi[kPiNumSlots] = nslot;
  pi[kPiOffHeader] = (int64_t)off_header;
  pi[kPiNumItems] = num_items;
  pi[kPiOffSegments] = (int64_t)off_segs;
  pi[kPiOffMerge] = (int64_t)off_merge;
  pi[kPiOffTmpO] = (int64_t)off_tmp_o;
  pi[kPiOffTmpLse] = (int64_t)off_tmp_lse;
  pi[kPiBatchSize] = batch_size;
  pi[kPiNumQoHeads] = num_qo_heads;
  pi[kPiNumKvHeads] = num_kv_heads;
  pi[kPiHeadDim] = head_dim;
  pi[kPiPageSize] = page_size;
  pi[kPiQDtype] = q_dtype;
  pi[kPiKvDtype] = kv_dtype;
  pi[kPiCudaGraph] = cuda_graph;
  // fused: the last CTA to arrive at a unit merges it inside the decode kernel; otherwise a
  // second launch (graph mode: fixed upper-bound grid so that replays stay valid)
  pi[kPiMergeGrid] = fused_merge ? 0 : (cuda_graph ? max_merge : nmerge);
  pi[kPiOffCounters] = (int64_t)off_counters;
  pi[kPiOffItemHeads] = (int64_t)off_heads;
  pi[kPiOffWorkCounter] = (int64_t)off_work;
  pi[kPiKvHeadDiv] = kv_div;
  pi[kPiTotalTiles] = total_tiles;
  pi[kPiWindowLeft] = window_left;
  pi[kPiStages] = cfg.stages;
  pi[kPiIntBytesUsed] = (int64_t)used;
  pi[kPiOffIndices] = idx_in_ws ? (int64_t)off_idx : -1;
  pi[kPiNumIndices] = idx_in_ws ? num_indices : 0;
  // the page table was fixed by host copies ordered before every later kernel of the stream (ours
  // above, or the wrapper's copy into the caller's CUDA-graph buffer): the decode kernel may read it
  // ahead of its programmatic-dependent-launch wait.  Tables the caller keeps on the device
  // (fast_decode_plan, device tensors handed to plan()) may be rewritten by an upstream kernel.
  pi[kPiIndicesStable] = kv_indices_host != nullptr ? 1 : 0;
  pi[kPiHeadPairs] = head_pairs ? 1 : 0;
  return FI_OK;
}

int fi_batch_decode_plan(void* float_ws, size_t float_ws_bytes, void* int_ws, void* pinned_int_ws,
                         size_t int_ws_bytes, const int32_t* indptr_host,
                         const int32_t* last_page_len_host, const int32_t* kv_len_host,
                         int batch_size, int num_qo_heads, int num_kv_heads, int head_dim,
                         int page_size, int q_dtype, int kv_dtype, int window_left,
                         int cuda_graph, int fixed_split_pages, int disable_split_kv, int device,
                         int num_ctas_override, void* stream, const int32_t* kv_indices_host,
                         int64_t num_indices, int64_t* plan_info_out) {
  return plan_impl(float_ws, float_ws_bytes, int_ws, pinned_int_ws, int_ws_bytes, indptr_host,
                   last_page_len_host, kv_len_host, batch_size, num_qo_heads, num_kv_heads, head_dim,
                   page_size, q_dtype, kv_dtype, window_left, cuda_graph, fixed_split_pages,
                   disable_split_kv, device, num_ctas_override, stream, kv_indices_host, num_indices,
                   plan_info_out, 0);
}

int fi_batch_decode_plan_args(const FiDecodePlanArgs* a) {
  if (a == nullptr) return fail(FI_ERR_INVALID_ARGUMENT, "null plan argument block");
  return plan_impl(a->float_ws, a->float_ws_bytes, a->int_ws, a->pinned_int_ws, a->int_ws_bytes,
                   a->indptr_host, a->last_page_len_host, a->kv_len_host, a->batch_size,
                   a->num_qo_heads, a->num_kv_heads, a->head_dim, a->page_size, a->q_dtype,
                   a->kv_dtype, a->window_left, a->cuda_graph, a->fixed_split_pages,
                   a->disable_split_kv, a->device, a->num_ctas_override, a->stream,
                   a->kv_indices_host, a->num_indices, a->plan_info_out, a->flags);
}

}  // extern "C"
