// decode_params.h — work-list records and kernel parameter block shared by the
// host planner (plan.cu), the launchers (decode.cu) and the kernels.
//
// Reference counterparts: DecodePlanInfo (include/flashinfer/attention/scheduler.cuh:366-422),
// Params (csrc/batch_decode_customize_config.jinja:28-60), paged_kv_t
// (include/flashinfer/page.cuh:37-210).  The design differs on purpose: instead of a
// (request, uniform-chunk) grid the planner cuts the flattened (request, kv-head, tile)
// stream into contiguous work items ("stream-K" over KV tiles) of decreasing size; each
// persistent CTA starts on item blockIdx.x and draws further items from an atomic counter,
// so SMs that stream faster take more of the tail (guided self-scheduling).
#pragma once
#include <stdint.h>

namespace fib200 {

constexpr int kConsumerWarps = 4;
// KV tokens per pipeline stage: 16 per consumer warp for 16-bit KV (one m16n8k16 M tile),
// 32 per warp for fp8 KV (one m16n8k32 K block of the P.V product).  A stage is 32 KiB at
// D = 128 either way.  kvf: 0 = 16-bit KV, 1 = e4m3, 2 = e5m2.
__host__ __device__ constexpr int tile_tokens(int kvf) { return kvf == 0 ? 64 : 128; }
constexpr int kMaxStages = 12;
constexpr int kScratchMlBytes = 768;   // per-CTA scratch (floats 0 .. 191): warp exchange, queue control, inline pages
constexpr int kTc5PairScratchBytes = 512;  // head-pair tcgen05 kernel only, after the barriers
// tcgen05 kernel modes: one KV head and <= 8 query columns per unit; a PAIR of KV heads (FI_PLAN_HEAD_PAIRS);
// one KV head and <= 16 (row, head) columns (FI_PLAN_WIDE_UNITS, FiDecodeRunArgs.q_rows)
constexpr int kTc5Single = 0, kTc5Pairs = 1, kTc5Wide = 2;
constexpr int kSegCache = 32;         // segment queue (ring) in smem per CTA; a power of two
constexpr int kSegRefill = 8;         // segments the feeder loads from the list at a time
constexpr float kNegInf = -5e4f;      // finite "-inf", reference math.cuh:32
constexpr int kMaxGroup = 8;          // GQA group size handled by one MMA N tile
// query heads one unit may carry: kMaxGroup, 4 at head_dim 512 (the CUDA-core variant kernel keeps the
// group's accumulators in registers)
__host__ __device__ constexpr int max_group_for(int head_dim) { return head_dim == 512 ? 4 : kMaxGroup; }
// smallest d with group % d == 0 and group / d <= max_group (1 for group <= max_group)
__host__ __device__ constexpr int group_split(int group, int max_group = kMaxGroup) {
  for (int d = 1; d <= group; ++d)
    if (group % d == 0 && group / d <= max_group) return d;
  return group;
}

// One unit of work for a CTA: tiles [tile_begin, tile_end) (absolute tile index on the
// request's token axis, tile t covers tokens [TT t, TT t + TT), TT = tile_tokens(kvf)) of KV head `kv_head`
// of request `request`.  Only tokens in [tok_lo, tok_hi) are attended.
// slot < 0: the segment covers the whole unit, write o/lse directly;
// slot >= 0: write a normalised partial (o fp32, lse) into partial slot `slot`.
struct alignas(16) Segment {
  int32_t request;
  int32_t kv_head;
  int32_t tile_begin;
  int32_t tile_end;
  int32_t slot;
  int32_t tok_lo;
  int32_t tok_hi;
  int32_t merge_unit;  // index of the MergeUnit this partial belongs to (-1 when slot < 0)
  int32_t page_base;   // kv_indptr[request]: first entry of the request in the page table
  int32_t merge_count; // copy of MergeUnit::count / slot_begin of `merge_unit`, so that the
  int32_t merge_slot;  //   fused merge never has to load the MergeUnit record
  int32_t kv_len;      // the request's full KV length (tok_hi may be a chunk end): RoPE q position
};
static_assert(sizeof(Segment) == 48, "Segment must be 48 bytes");

// Head of one work item: the item's segment range plus its first kItemInlineSegs segments
// inline and — when plan() was given the page table on the host — the page ids of the item's
// first tiles, so that one 224-byte read (the kernel prologue for item blockIdx.x, the feeder for
// drawn items) is all that stands between a CTA and its first TMA box: no page-table lookup on the
// critical path of the launch.  Segments beyond the inline ones are read from the list, which
// starts right after the last head.
constexpr int kItemInlineSegs = 3;
constexpr int kItemInlinePages = 16;
struct alignas(16) ItemHead {
  int32_t first;      // index of the item's first segment in the list
  int32_t count;      // segments in the item
  int32_t num_items;  // items in the plan (same in every head)
  int32_t num_pages;  // valid entries of pages[] (0: look the pages up in the table)
  Segment seg[kItemInlineSegs];
  // kv_indices[seg[0].page_base + first_page + i], first_page = page of token
  // max(seg[0].tile_begin * tile_tokens, 0): the pages the item's first tiles gather
  int32_t pages[kItemInlinePages];
};
static_assert(sizeof(ItemHead) == 224, "ItemHead must be 224 bytes");

// A (request, kv_head) unit whose KV range was split over several CTAs: partial
// slots [slot_begin, slot_begin + count) are merged into o[request, kv_head*G ..].
struct alignas(16) MergeUnit {
  int32_t request;
  int32_t kv_head;
  int32_t slot_begin;
  int32_t count;
};

// Device-resident header (first bytes of the int workspace), rewritten by every plan()
// so that CUDA-graph replays pick up new counts without new launch geometry.
struct PlanHeader {
  int32_t num_items;
  int32_t num_segments;
  int32_t num_merge_units;
  int32_t num_slots;
  int32_t total_tiles;
  int32_t pad[3];
};

// indices into the int64 plan-info vector handed to Python (opaque there)
enum PlanInfoSlot {
  kPiMagic = 0,
  kPiNumCtas,       // launch grid of the decode kernel
  kPiNumSegments,
  kPiNumMergeUnits,
  kPiNumSlots,
  kPiOffHeader,
  kPiNumItems,
  kPiOffSegments,
  kPiOffMerge,
  kPiOffTmpO,
  kPiOffTmpLse,
  kPiBatchSize,
  kPiNumQoHeads,
  kPiNumKvHeads,
  kPiHeadDim,
  kPiPageSize,
  kPiQDtype,
  kPiKvDtype,
  kPiCudaGraph,
  kPiMergeGrid,     // launch grid of the merge kernel (0: no merge launch)
  kPiTotalTiles,
  kPiWindowLeft,
  kPiStages,
  kPiIntBytesUsed,
  kPiOffCounters,   // int32 arrival counters, one per merge unit (fused split-KV merge)
  kPiOffItemHeads,
  kPiOffWorkCounter,  // int32 ticket counter of the item scheduler (zero between launches)
  kPiKvHeadDiv,       // virtual KV heads per real one (kPiNumKvHeads counts virtual heads)
  kPiOffIndices,      // byte offset in the int workspace of plan()'s device copy of kv_indices, or -1
  kPiNumIndices,      // entries of that copy
  kPiIndicesStable,   // 1: kv_indices is written by host copies only (plan() had it on the host)
  kPiHeadPairs,       // 1: units are (request, PAIR of adjacent KV heads), tiles of 64 tokens (fp8 tcgen05 kernel)
  kPiLen
};
constexpr int64_t kPlanMagic = 0xF1B2000001LL;

struct DecodeParams {
  // q / o
  const void* q;
  int64_t q_stride_n, q_stride_h;  // elements
  void* o;                         // [bs, Hq, D] contiguous, q dtype
  float* lse;                      // [bs, Hq] or null
  // kv (element strides; used by the cp.async loader, TMA uses the tensor maps)
  const void* k;
  const void* v;
  int64_t kv_stride_page, kv_stride_n, kv_stride_h;
  const int32_t* kv_indices;  // null: identity page table
  const int32_t* kv_indptr;   // null: page base 0 (contiguous single request)
  int32_t page_size;
  int32_t page_shift;  // log2(page_size) when a power of two, else -1 (31: one giant page)
  int32_t page_mask;   // page_size - 1 when a power of two (0x7fffffff: one giant page)
  int32_t rows_per_box;  // TMA: min(page_size, 64)
  int32_t tma_h_first;   // TMA map dim order: 1 = (D, h, n, page), 0 = (D, n, h, page)
  int32_t tma_paired;    // 1: kmap is 5-D (D, n|h, h|n, K|V, page): one box brings K and V of a token block
  // work list (explicit) ...
  const PlanHeader* header;        // non-null selects this mode (the decode kernel never reads it)
  const ItemHead* item_heads;      // the segment list follows head num_items - 1
  int32_t* work_counter;           // tickets: item = ticket + gridDim.x; returns to zero
  int32_t sched_lookahead;         // draw the next item when <= this many published tiles are left
  // ... or implicit uniform mode (single request, contiguous KV): segments are
  // derived in closed form from these (header == nullptr).
  int32_t imp_units;      // number of kv heads
  int32_t imp_tiles;      // tiles per unit
  int32_t imp_tok_lo, imp_tok_hi;
  // partial outputs
  float* tmp_o;    // [slots, G, D]
  float* tmp_lse;  // [slots, G]
  // fused split-KV merge (explicit work list only): the CTA whose partial arrives last at a
  // unit merges all of the unit's partials; counters return to zero for the next run()
  const MergeUnit* merge_units;
  int32_t* merge_counters;
  int32_t num_qo_heads;
  int32_t group_size;
  int32_t num_stages;
  float sm_scale_log2;     // sm_scale * log2(e)
  float v_scale;
  float soft_cap;          // <= 0 disabled
  const float* alibi_slopes;  // [num_qo_heads] or null: logit += slope[head] * kv position (ALiBi)
  int32_t rope;               // 1: rotate q (at kv_len - 1) and every K row (at its position), LLaMA style
  float rope_rcp_scale, rope_rcp_theta;
  float sm_scale;          // plain sm_scale (soft-cap path)
  int32_t enable_pdl;
  int32_t indices_early;  // 1: kv_indices may be read before the PDL dependency wait (plan-owned copy)
  int32_t l2_hints;  // bit 0: KV boxes L2 evict_first; bit 1: page table / work list L2 evict_last
  // GQA groups larger than kMaxGroup are run as `kv_head_div` virtual KV heads per real one, each
  // with group / kv_head_div (<= kMaxGroup) query heads: Segment::kv_head counts virtual heads
  // (query / output addressing), the cache is addressed with kv_head / kv_head_div
  int32_t kv_head_div;
  // head-pair plan (fp8 KV on the tcgen05 kernel): Segment::kv_head counts PAIRS of adjacent KV heads,
  // group_size is twice the real GQA group (the query heads of both), tiles are 64 tokens
  int32_t head_pairs;
  // multi-row units (run(q_len_per_req = n), 16-bit KV): q / o hold q_rows consecutive rows per request
  // and num_qo_heads = q_rows x the model's query heads; q_group = the model's GQA group; the unit's
  // group_size columns are (row, head) pairs (decode_kernel.cuh row_head).  q_rows <= 1: off
  int32_t q_rows, q_group;
  int32_t q_rows_noncausal;  // 1: no causal mask among the rows (cascade attention's shared levels)
  // attention sinks (reference decode.py:1176, tests/test_helpers/sink_attention_reference.py:23-36):
  // one extra logit per query head that takes part in the softmax denominator only —
  // p_i = exp(s_i) / (sum_j exp(s_j) + exp(sink[head])); natural-log units, not scaled by sm_scale.
  // Added once per (request, head): where the final o / lse are written, never to a partial.
  const float* sinks;
};

struct MergeParams {
  const PlanHeader* header;     // null in implicit mode
  const MergeUnit* units;
  int32_t imp_units, imp_tiles, imp_ctas;  // implicit mode
  const float* tmp_o;
  const float* tmp_lse;
  void* o;
  float* lse;
  int32_t num_qo_heads, group_size, head_dim;
  float out_scale;
  int32_t out_dtype;
  int32_t enable_pdl;
  const float* sinks;  // see DecodeParams::sinks
};

}  // namespace fib200
