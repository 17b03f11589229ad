// decode.cu — launchers and C-ABI entry points for batch / single decode.
// Replaces BatchDecodeWithPagedKVCacheRun (reference csrc/batch_decode.cu:81-191),
// BatchDecodeWithPagedKVCacheDispatched (include/flashinfer/attention/decode.cuh:739-825),
// single_decode_with_kv_cache (csrc/single_decode.cu:33-104) and
// SingleDecodeWithKVCacheDispatched (decode.cuh:658-737).
#include <cuda.h>
#include <cudaTypedefs.h>
#include <stdlib.h>
#include <string.h>

#include <map>
#include <mutex>
#include <utility>

#ifdef FI_B200_TRACE
// the trace build (build.py --trace) is one translation unit: the kernels share g_fi_trace
#include "decode_mma.cu"
#include "decode_tc5.cu"
#include "decode_variant.cu"
#endif
#include "decode_launch.h"
#include "host_common.h"
#include "merge_kernels.cuh"

namespace fib200 {

// ---- TMA descriptor ------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
            cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

// 4-D map over one of K / V: dims (D, n, h, page) — or (D, h, n, page) when the head
// stride is the smaller one (NHD), so that strides grow with the dimension index.  With
// stride_kv > 0 the map is 5-D, (D, n|h, h|n, K|V, page), over K and V together: one box
// {128 B, rows_per_box, 1, 2, 1} brings the K and the V rows of a token block.
// A descriptor is a pure function of these arguments; a serving loop calls run() with the same
// few caches (one per layer), so the last encodings are kept per thread.
constexpr int64_t kHeadPairMap = -1;  // stride_kv value that selects the head-pair map

struct KvMapKey {
  const void* base;
  int64_t dim_n, dim_h, dim_page, stride_n, stride_h, stride_page, stride_kv;
  int kv_dtype, head_dim, rows_per_box;
  bool operator==(const KvMapKey& o) const {
    return base == o.base && dim_n == o.dim_n && dim_h == o.dim_h && dim_page == o.dim_page &&
           stride_n == o.stride_n && stride_h == o.stride_h && stride_page == o.stride_page &&
           stride_kv == o.stride_kv && kv_dtype == o.kv_dtype && head_dim == o.head_dim &&
           rows_per_box == o.rows_per_box;
  }
};
constexpr int kMapCacheSize = 256;  // K and V of 128 layers
struct KvMapCache {
  KvMapKey key[kMapCacheSize];
  CUtensorMap map[kMapCacheSize];
  int used = 0, next = 0;
};

static int encode_kv_map(CUtensorMap* map, const void* base, int kv_dtype, int head_dim,
                         int64_t dim_n, int64_t dim_h, int64_t dim_page, int64_t stride_n,
                         int64_t stride_h, int64_t stride_page, int rows_per_box, int64_t stride_kv);

static int make_kv_map(CUtensorMap* map, const void* base, int kv_dtype, int head_dim,
                       int64_t dim_n, int64_t dim_h, int64_t dim_page, int64_t stride_n,
                       int64_t stride_h, int64_t stride_page, int rows_per_box, int64_t stride_kv = 0) {
  static thread_local KvMapCache cache;
  const KvMapKey key = {base, dim_n, dim_h, dim_page, stride_n, stride_h, stride_page, stride_kv,
                        kv_dtype, head_dim, rows_per_box};
  for (int i = 0; i < cache.used; ++i)
    if (cache.key[i] == key) { *map = cache.map[i]; return FI_OK; }
  const int rc = encode_kv_map(map, base, kv_dtype, head_dim, dim_n, dim_h, dim_page, stride_n,
                               stride_h, stride_page, rows_per_box, stride_kv);
  if (rc != FI_OK) return rc;
  const int slot = cache.used < kMapCacheSize ? cache.used++ : (cache.next++ % kMapCacheSize);
  cache.key[slot] = key;
  cache.map[slot] = *map;
  return FI_OK;
}

static int encode_kv_map(CUtensorMap* map, const void* base, int kv_dtype, int head_dim,
                         int64_t dim_n, int64_t dim_h, int64_t dim_page, int64_t stride_n,
                         int64_t stride_h, int64_t stride_page, int rows_per_box, int64_t stride_kv) {
  EncodeTiledFn enc = get_encode_fn();
  if (enc == nullptr) return fail(FI_ERR_CUDA, "cuTensorMapEncodeTiled entry point not found");
  const int es = dtype_size(kv_dtype);
  CUtensorMapDataType dt = es == 1 ? CU_TENSOR_MAP_DATA_TYPE_UINT8
                                   : (kv_dtype == FI_DTYPE_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16
                                                                : CU_TENSOR_MAP_DATA_TYPE_FLOAT16);
  if (stride_kv == kHeadPairMap) {
    // (128 B, token % 8, head, token / 8, page), box {128, 8, 2, rows_per_box / 8, 1}: the box lands as
    // 8-row groups ordered (token / 8, head), 1 KiB each = one 128B-swizzle atom (decode_kernel.cuh
    // produce_tma HEADPAIR).  fp8 KV at head_dim 128 only.
    cuuint64_t d5[5] = {128, 8, (cuuint64_t)dim_h, (cuuint64_t)(dim_n / 8), (cuuint64_t)dim_page};
    cuuint64_t s4[4] = {(cuuint64_t)stride_n, (cuuint64_t)stride_h, (cuuint64_t)(8 * stride_n),
                        (cuuint64_t)stride_page};
    cuuint32_t b5[5] = {128, 8, 2, (cuuint32_t)(rows_per_box / 8), 1};
    cuuint32_t e5[5] = {1, 1, 1, 1, 1};
    // the pair's two rows are adjacent in an NHD cache: fetch them as one 256-byte request
    const CUtensorMapL2promotion promo5 = stride_h == 128 ? CU_TENSOR_MAP_L2_PROMOTION_L2_256B
                                                          : CU_TENSOR_MAP_L2_PROMOTION_L2_128B;
    CUresult r5 = enc(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 5, const_cast<void*>(base), d5, s4, b5, e5,
                      CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, promo5,
                      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r5 != CUDA_SUCCESS) return fail(FI_ERR_CUDA, "cuTensorMapEncodeTiled (head pairs) failed with %d", (int)r5);
    return FI_OK;
  }
  const bool five = stride_kv > 0;
  cuuint64_t dims[5] = {(cuuint64_t)head_dim, (cuuint64_t)dim_n, (cuuint64_t)dim_h,
                        (cuuint64_t)dim_page, 1};
  cuuint64_t strides[4] = {(cuuint64_t)(stride_n * es), (cuuint64_t)(stride_h * es),
                           (cuuint64_t)(stride_page * es), 0};
  cuuint32_t box[5] = {(cuuint32_t)(128 / es), (cuuint32_t)rows_per_box, 1, 1, 1};
  if (five) {  // (D, n, h, K|V, page)
    dims[3] = 2; dims[4] = (cuuint64_t)dim_page;
    strides[2] = (cuuint64_t)(stride_kv * es); strides[3] = (cuuint64_t)(stride_page * es);
    box[3] = 2;
  }
  if (stride_h < stride_n) {
    std::swap(dims[1], dims[2]);
    std::swap(strides[0], strides[1]);
    std::swap(box[1], box[2]);
  }
  cuuint32_t estr[5] = {1, 1, 1, 1, 1};
  // L2 promotion must not exceed the contiguous run of one (token, head) row: a 128-byte row
  // (fp8 D=128, 16-bit D=64) inside an NHD page has another head's row next to it, and
  // promoting to 256 B would drag that row through DRAM for nothing (measured 1.49x traffic).
  const int64_t row_bytes = (int64_t)head_dim * es;
  const bool rows_adjacent = stride_n * es == row_bytes;
  const CUtensorMapL2promotion promo = (row_bytes >= 256 || rows_adjacent)
                                           ? CU_TENSOR_MAP_L2_PROMOTION_L2_256B
                                           : CU_TENSOR_MAP_L2_PROMOTION_L2_128B;
  CUresult r = enc(map, dt, five ? 5 : 4, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, promo,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(FI_ERR_CUDA, "cuTensorMapEncodeTiled failed with %d", (int)r);
  return FI_OK;
}

static bool is_pow2(int64_t x) { return x > 0 && (x & (x - 1)) == 0; }
static int ilog2(int64_t x) { int s = 0; while ((int64_t(1) << s) < x) ++s; return s; }

static bool tma_eligible(const void* k, const void* v, int es, int64_t sn, int64_t sh, int64_t sp,
                         int page_size, bool contiguous, int64_t kv_len, int tile) {
  if (tuning().loader == 2) return false;
  if ((reinterpret_cast<uintptr_t>(k) | reinterpret_cast<uintptr_t>(v)) & 15) return false;
  if ((sn * es) % 16 || (sh * es) % 16 || (sp * es) % 16) return false;
  if (contiguous) return kv_len >= tile;
  return is_pow2(page_size) && page_size >= 8;
}

// ---- kernel launch ---------------------------------------------------------------
// The kernels are instantiated in decode_mma.cu / decode_tc5.cu / decode_variant.cu (decode_launch.h).
static int default_l2_hints() {
  static const int l2_env = [] { const char* e = getenv("FI_B200_L2_HINTS"); return e ? atoi(e) : -1; }();
  return l2_env >= 0 ? l2_env : 3;
}

static bool use_tc5() {
  static const bool on = [] {
    const char* e = getenv("FI_B200_FP8_TC5");
    return !(e != nullptr && e[0] == '0');
  }();
  return on;
}

static int launch_decode(const DecodeParams& p, const CUtensorMap& km, const CUtensorMap& vm,
                         int q_dtype, int kvf, int head_dim, int grid, bool tma,
                         cudaStream_t stream) {
  const bool bf = q_dtype == FI_DTYPE_BF16;
  static const bool rope_cuda_core = [] { const char* e = getenv("FI_B200_ROPE_CUDA_CORE"); return e && e[0] == '1'; }();
  // fp8 KV with ROPE_LLAMA (or A/B runs), and fp8 KV at head_dim 64 in any mode but ALiBi (64-byte rows do
  // not fit the 128-byte swizzle atoms of the MMA kernels): the CUDA-core variant kernel
  // (decode_rope_kernel.cuh); without RoPE it runs with zero rotation frequencies, i.e. the identity rotation
  if ((p.rope && (kvf != 0 || rope_cuda_core)) || (kvf != 0 && head_dim == 64) || head_dim == 512) {
    if (p.alibi_slopes != nullptr)
      return fail(FI_ERR_UNSUPPORTED, "ALIBI at head_dim 512 or with fp8 KV at head_dim 64 is not built");
    DecodeParams p2 = p;
    if (!p.rope) {
      p2.rope_rcp_scale = 0.f;
      p2.rope_rcp_theta = 1.f;
    }
    return bf ? launch_variant_slice<1>(p2, kvf, head_dim, grid, stream)
              : launch_variant_slice<0>(p2, kvf, head_dim, grid, stream);
  }
  // fp8 KV on the tcgen05 / TMEM kernels (decode_kernel_tc5.cuh)
  int tc5_mode = -1;
  if (p.head_pairs) tc5_mode = kTc5Pairs;  // fi_batch_decode_run checked that this launch fits a head-pair plan
  else if (kvf != 0 && p.q_rows >= 1) tc5_mode = kTc5Wide;  // wide / multi-row units (checked by the caller: D 128, TMA, tcgen05)
  else if (kvf != 0 && (head_dim == 256 || head_dim == 128) && tma && use_tc5()) tc5_mode = kTc5Single;
  if (tc5_mode >= 0)
    return bf ? launch_tc5_slice<1>(p, km, vm, kvf, tc5_mode, head_dim, grid, stream)
              : launch_tc5_slice<0>(p, km, vm, kvf, tc5_mode, head_dim, grid, stream);
  if (kvf == 0)
    return bf ? launch_mma_slice<1, false>(p, km, vm, kvf, head_dim, grid, tma, stream)
              : launch_mma_slice<0, false>(p, km, vm, kvf, head_dim, grid, tma, stream);
  return bf ? launch_mma_slice<1, true>(p, km, vm, kvf, head_dim, grid, tma, stream)
            : launch_mma_slice<0, true>(p, km, vm, kvf, head_dim, grid, tma, stream);
}

static int launch_merge(const MergeParams& mp, int grid, cudaStream_t stream) {
  if (grid <= 0) return FI_OK;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(128);
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = mp.enable_pdl ? 1 : 0;
  if (mp.out_dtype == FI_DTYPE_BF16) FI_CUDA_TRY(cudaLaunchKernelEx(&cfg, decode_merge_kernel<1>, mp));
  else FI_CUDA_TRY(cudaLaunchKernelEx(&cfg, decode_merge_kernel<0>, mp));
  return FI_OK;
}

static int check_pos_mode(int mode, const float* alibi_slopes, float rope_scale, float rope_theta) {
  if (mode != FI_POS_NONE && mode != FI_POS_ROPE_LLAMA && mode != FI_POS_ALIBI)
    return fail(FI_ERR_INVALID_ARGUMENT, "invalid pos_encoding_mode %d", mode);
  if (mode == FI_POS_ALIBI && alibi_slopes == nullptr)
    return fail(FI_ERR_INVALID_ARGUMENT, "ALIBI needs alibi_slopes");
  if (mode == FI_POS_ROPE_LLAMA && !(rope_scale > 0.f && rope_theta > 0.f))
    return fail(FI_ERR_INVALID_ARGUMENT, "ROPE_LLAMA needs rope_scale > 0 and rope_theta > 0");
  return FI_OK;
}

static int g_last_loader = -1;  // 0 TMA, 1 cp.async: what the last run() launched (for tests)
static long long g_launch_count = 0;

}  // namespace fib200

using namespace fib200;

extern "C" {

#ifdef FI_B200_TRACE
// trace build only: copy the per-CTA phase timestamps (kTraceSlots per CTA) to the host
FI_API int fi_debug_trace(unsigned long long* dst, int n_ctas) {
  if (n_ctas > kTraceCtas) n_ctas = kTraceCtas;
  cudaError_t e = cudaMemcpyFromSymbol(dst, g_fi_trace, sizeof(unsigned long long) * kTraceSlots * n_ctas);
  if (e != cudaSuccess) return fail(FI_ERR_CUDA, "trace copy failed: %s", cudaGetErrorString(e));
  unsigned long long* sym = nullptr;
  if (cudaGetSymbolAddress(reinterpret_cast<void**>(&sym), g_fi_trace) == cudaSuccess)
    cudaMemset(sym, 0, sizeof(unsigned long long) * kTraceSlots * kTraceCtas);
  return FI_OK;
}
#endif

int fi_last_loader(void) { return g_last_loader; }
long long fi_launch_count(void) { return g_launch_count; }

static int batch_decode_run_impl(void* float_ws, void* int_ws, const int64_t* plan_info, const void* q,
                                 int64_t q_stride_n, int64_t q_stride_h, const void* k_cache,
                                 const void* v_cache, int64_t num_pages, int64_t kv_stride_page,
                                 int64_t kv_stride_n, int64_t kv_stride_h, const int32_t* kv_indices,
                                 const int32_t* kv_indptr, void* o, float* lse, float sm_scale,
                                 float v_scale, float logits_soft_cap, int pos_encoding_mode,
                                 const float* alibi_slopes, float rope_scale, float rope_theta,
                                 const float* sinks, int enable_pdl, int device, void* stream_, int q_rows,
                                 int q_rows_noncausal = 0) {
  if (int rc = check_pos_mode(pos_encoding_mode, alibi_slopes, rope_scale, rope_theta)) return rc;
  if (plan_info == nullptr || plan_info[kPiMagic] != kPlanMagic)
    return fail(FI_ERR_INVALID_ARGUMENT, "plan info is not initialized");
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const int bs = (int)plan_info[kPiBatchSize];
  if (bs == 0) return FI_OK;
  const int q_dtype = (int)plan_info[kPiQDtype], kv_dtype = (int)plan_info[kPiKvDtype];
  const int kvf = kv_format(q_dtype, kv_dtype);
  const int es = dtype_size(kv_dtype);
  const int head_dim = (int)plan_info[kPiHeadDim];
  const int page_size = (int)plan_info[kPiPageSize];
  const int Hq = (int)plan_info[kPiNumQoHeads], Hkv = (int)plan_info[kPiNumKvHeads];
  if ((q_stride_n % 4) || (q_stride_h % 4) || (reinterpret_cast<uintptr_t>(q) & 7))
    return fail(FI_ERR_INVALID_ARGUMENT, "q rows must be 8-byte aligned");
  if ((kv_stride_n * es) % 16 || (kv_stride_h * es) % 16 || (kv_stride_page * es) % 16 ||
      ((reinterpret_cast<uintptr_t>(k_cache) | reinterpret_cast<uintptr_t>(v_cache)) & 15))
    return fail(FI_ERR_INVALID_ARGUMENT, "k/v cache rows must be 16-byte aligned");
  FI_CUDA_TRY(cudaSetDevice(device));

  char* iw = static_cast<char*>(int_ws);
  char* fw = static_cast<char*>(float_ws);
  DecodeParams p;
  memset(&p, 0, sizeof(p));
  p.q = q; p.q_stride_n = q_stride_n; p.q_stride_h = q_stride_h;
  p.o = o; p.lse = lse;
  p.k = k_cache; p.v = v_cache;
  p.kv_stride_page = kv_stride_page; p.kv_stride_n = kv_stride_n; p.kv_stride_h = kv_stride_h;
  p.kv_indices = kv_indices; p.kv_indptr = kv_indptr;
  // no page table passed: plan() put its own device copy behind the work list (it had the table on
  // the host).  NULL without such a copy keeps its old meaning: identity page table.
  if (kv_indices == nullptr && plan_info[kPiOffIndices] >= 0)
    p.kv_indices = reinterpret_cast<const int32_t*>(iw + plan_info[kPiOffIndices]);
  p.indices_early = plan_info[kPiIndicesStable] != 0 || p.kv_indices == nullptr;
  p.page_size = page_size;
  p.page_shift = is_pow2(page_size) ? ilog2(page_size) : -1;
  p.page_mask = is_pow2(page_size) ? page_size - 1 : 0;
  p.head_pairs = plan_info[kPiHeadPairs] != 0;
  const int TT = p.head_pairs ? tile_tokens(kvf) / 2 : tile_tokens(kvf);
  p.rows_per_box = page_size < TT ? page_size : TT;
  p.tma_h_first = kv_stride_h < kv_stride_n;
  p.header = reinterpret_cast<const PlanHeader*>(iw + plan_info[kPiOffHeader]);
  p.item_heads = reinterpret_cast<const ItemHead*>(iw + plan_info[kPiOffItemHeads]);
  p.work_counter = reinterpret_cast<int32_t*>(iw + plan_info[kPiOffWorkCounter]);
  p.sched_lookahead = sched_tuning().lookahead_tiles;
  p.tmp_o = reinterpret_cast<float*>(fw + plan_info[kPiOffTmpO]);
  p.tmp_lse = reinterpret_cast<float*>(fw + plan_info[kPiOffTmpLse]);
  p.merge_units = reinterpret_cast<const MergeUnit*>(iw + plan_info[kPiOffMerge]);
  p.merge_counters = reinterpret_cast<int32_t*>(iw + plan_info[kPiOffCounters]);
  p.num_qo_heads = Hq;
  p.group_size = Hq / Hkv;  // Hkv counts virtual heads: <= kMaxGroup query heads each
  p.kv_head_div = std::max<int>((int)plan_info[kPiKvHeadDiv], 1);
  p.num_stages = (int)plan_info[kPiStages];
  p.sm_scale = sm_scale;
  p.sm_scale_log2 = sm_scale * 1.4426950408889634f;
  p.v_scale = v_scale;
  p.soft_cap = logits_soft_cap;
  p.alibi_slopes = pos_encoding_mode == FI_POS_ALIBI ? alibi_slopes : nullptr;
  p.rope = pos_encoding_mode == FI_POS_ROPE_LLAMA;
  p.rope_rcp_scale = 1.f / rope_scale; p.rope_rcp_theta = 1.f / rope_theta;
  if (((p.rope && kvf != 0) || (kvf != 0 && head_dim == 64) || head_dim == 512) &&
      plan_info[kPiNumItems] > plan_info[kPiNumCtas])
    return fail(FI_ERR_UNSUPPORTED, "the CUDA-core variant kernel (fp8 KV with ROPE_LLAMA or head_dim 64) needs a "
                                    "static plan (unset FI_B200_SCHED_F0)");
  // 1: griddepcontrol.launch_dependents right after the kernel's own dependency wait (default);
  // 2: at the very end (FI_B200_PDL_LATE=1, kept for A/B runs)
  static const bool pdl_late = [] { const char* e = getenv("FI_B200_PDL_LATE"); return e && e[0] == '1'; }();
  p.enable_pdl = enable_pdl ? (pdl_late ? 2 : 1) : 0;
  // L2 eviction hints on the TMA path (FI_B200_L2_HINTS overrides): bit 0 = the KV stream is read once
  // -> evict_first, bit 1 = page table / work list are re-read by every layer's launch -> evict_last.
  // Measured on B200 with both: C3 e4m3 368 -> 353 us, its 8-way per-rank shape 60.9 -> 54.7 us,
  // C2 165.4 -> 163.7 us, per-rank 29.4 -> 28.8 us; with another kernel's clean data in L2 at launch
  // (scripts/l2_context_probe.py) C3 385 -> 362 us, C2 167.9 -> 165.9 us.  Only a launch that finds
  // the L2 full of DIRTY lines (a 256 MB memset right before it) is slower (C2 +4 %, C4 +10 %).
  p.l2_hints = default_l2_hints();
  p.sinks = sinks;
  const bool wide_plan = !p.head_pairs && p.group_size > kMaxGroup;  // FI_PLAN_WIDE_UNITS took effect
  // fp8 KV: the tcgen05 kernel's wide mode (head_dim 128, TMA pages: checked below with `tma`)
  const bool fp8_wide_ok = kvf != 0 && head_dim == 128 && use_tc5() && !p.head_pairs;
  if (wide_plan && (p.rope || (kvf != 0 && !fp8_wide_ok)))
    return fail(FI_ERR_REPLAN, "the plan was made with FI_PLAN_WIDE_UNITS, but this run() needs %s",
                p.rope ? "ROPE_LLAMA" : "the register path");
  if (wide_plan && q_rows <= 1) {
    // a GQA group of 9 .. 16 as ONE unit: the multi-row consumer with a single row per request
    p.q_rows = 1;
    p.q_group = Hq / (Hkv / p.kv_head_div);
  }
  if (q_rows > 1) {
    // multi-row units: the plan was made with num_qo_heads = q_rows x the model's query heads
    const int Hkv_mem = Hkv / p.kv_head_div;
    if ((kvf != 0 && !fp8_wide_ok) || p.rope || Hq % q_rows != 0 || (Hq / q_rows) % Hkv_mem != 0 || p.head_pairs ||
        plan_info[kPiMergeGrid] > 0 || plan_info[kPiWindowLeft] >= 0 || head_dim == 512)
      return fail(FI_ERR_UNSUPPORTED, "q_rows > 1 needs 16-bit KV or fp8 KV at head_dim 128, no ROPE_LLAMA, no sliding "
                                      "window, the fused merge and a plan made with num_qo_heads = q_rows x query heads");
    p.q_rows = q_rows;
    p.q_rows_noncausal = q_rows_noncausal != 0;
    p.q_group = (Hq / q_rows) / Hkv_mem;
  }

  const bool cuda_core_only = (kvf != 0 && head_dim == 64) || head_dim == 512;
  const bool tma = !cuda_core_only && tma_eligible(k_cache, v_cache, es, kv_stride_n, kv_stride_h,
                                                   kv_stride_page, page_size, false, 0, TT);
  if (kvf != 0 && p.q_rows >= 1 && !tma)  // fp8 wide / multi-row units exist on the tcgen05 kernel only
    return fail(wide_plan ? FI_ERR_REPLAN : FI_ERR_UNSUPPORTED,
                "fp8 KV with wide or multi-row units needs TMA-addressable pages (a power of two >= 8)");
  if (p.head_pairs && (!tma || p.rope || !use_tc5() || es != 1 || head_dim != 128))
    return fail(FI_ERR_REPLAN, "the plan was made with FI_PLAN_HEAD_PAIRS, but this run() needs %s: plan "
                               "again without the flag", p.rope ? "ROPE_LLAMA" : "the gather loader");
  if (!tma && tuning().loader == 1)
    return fail(FI_ERR_UNSUPPORTED, "TMA loader forced but page geometry is not eligible");
  if (!tma && (kv_stride_page * es >= (int64_t(1) << 32) || kv_stride_n * es >= (int64_t(1) << 32)))
    return fail(FI_ERR_UNSUPPORTED, "page / token strides of 4 GiB or more are not supported by the gather loader");
  CUtensorMap km, vm;
  memset(&km, 0, sizeof(km));
  memset(&vm, 0, sizeof(vm));
  if (tma && p.head_pairs) {
    int rc = make_kv_map(&km, k_cache, kv_dtype, head_dim, page_size, 2 * Hkv, num_pages, kv_stride_n,
                         kv_stride_h, kv_stride_page, p.rows_per_box, kHeadPairMap);
    if (rc != FI_OK) return rc;
    rc = make_kv_map(&vm, v_cache, kv_dtype, head_dim, page_size, 2 * Hkv, num_pages, kv_stride_n,
                     kv_stride_h, kv_stride_page, p.rows_per_box, kHeadPairMap);
    if (rc != FI_OK) return rc;
  } else if (tma) {
    const int Hkv_mem = Hkv / p.kv_head_div;  // heads the cache really has
    // K and V of a page interleaved inside the page stride (the 5-D cache every engine allocates):
    // one 5-D map over both, one TMA box per (page, token block) instead of two (mma.sync kernel)
    const int64_t kv_gap = (static_cast<const char*>(v_cache) - static_cast<const char*>(k_cache)) / es;
    static const bool pair_env = [] { const char* e = getenv("FI_B200_TMA_PAIRED"); return !(e && e[0] == '0'); }();
    const bool tc5 = kvf != 0 && (head_dim == 128 || head_dim == 256) && use_tc5() && !p.rope;
    p.tma_paired = pair_env && !tc5 && v_cache > k_cache && kv_gap * es % 16 == 0 && kv_gap < kv_stride_page &&
                   (static_cast<const char*>(v_cache) - static_cast<const char*>(k_cache)) % es == 0;
    int rc = make_kv_map(&km, k_cache, kv_dtype, head_dim, page_size, Hkv_mem, num_pages, kv_stride_n,
                         kv_stride_h, kv_stride_page, p.rows_per_box, p.tma_paired ? kv_gap : 0);
    if (rc != FI_OK) return rc;
    if (!p.tma_paired) {
      rc = make_kv_map(&vm, v_cache, kv_dtype, head_dim, page_size, Hkv_mem, num_pages, kv_stride_n,
                       kv_stride_h, kv_stride_page, p.rows_per_box);
      if (rc != FI_OK) return rc;
    }
  }
  int rc = launch_decode(p, km, vm, q_dtype, kvf, head_dim, (int)plan_info[kPiNumCtas], tma, stream);
  if (rc != FI_OK) return rc;
  g_last_loader = tma ? 0 : 1;
  ++g_launch_count;

  const int merge_grid = (int)plan_info[kPiMergeGrid];
  if (merge_grid > 0) {
    MergeParams mp;
    memset(&mp, 0, sizeof(mp));
    mp.header = p.header;
    mp.units = reinterpret_cast<const MergeUnit*>(iw + plan_info[kPiOffMerge]);
    mp.tmp_o = p.tmp_o; mp.tmp_lse = p.tmp_lse;
    mp.o = o; mp.lse = lse;
    mp.num_qo_heads = Hq; mp.group_size = p.group_size; mp.head_dim = head_dim;
    mp.sinks = sinks;
    mp.out_scale = v_scale;
    mp.out_dtype = q_dtype;
    mp.enable_pdl = enable_pdl;
    rc = launch_merge(mp, merge_grid, stream);
    if (rc != FI_OK) return rc;
    ++g_launch_count;
  }
  return FI_OK;
}

int fi_batch_decode_run(void* float_ws, void* int_ws, const int64_t* plan_info, const void* q,
                        int64_t q_stride_n, int64_t q_stride_h, const void* k_cache,
                        const void* v_cache, int64_t num_pages, int64_t kv_stride_page,
                        int64_t kv_stride_n, int64_t kv_stride_h, const int32_t* kv_indices,
                        const int32_t* kv_indptr, void* o, float* lse, float sm_scale,
                        float v_scale, float logits_soft_cap, int pos_encoding_mode,
                        const float* alibi_slopes, float rope_scale, float rope_theta,
                        const float* sinks, int enable_pdl, int device, void* stream_) {
  return batch_decode_run_impl(float_ws, int_ws, plan_info, q, q_stride_n, q_stride_h, k_cache, v_cache,
                               num_pages, kv_stride_page, kv_stride_n, kv_stride_h, kv_indices, kv_indptr, o,
                               lse, sm_scale, v_scale, logits_soft_cap, pos_encoding_mode, alibi_slopes,
                               rope_scale, rope_theta, sinks, enable_pdl, device, stream_, 1);
}

int fi_batch_decode_run_args(const FiDecodeRunArgs* a) {
  if (a == nullptr) return fail(FI_ERR_INVALID_ARGUMENT, "null run argument block");
  return batch_decode_run_impl(a->float_ws, a->int_ws, a->plan_info, a->q, a->q_stride_n, a->q_stride_h,
                               a->k_cache, a->v_cache, a->num_pages, a->kv_stride_page, a->kv_stride_n,
                               a->kv_stride_h, a->kv_indices, a->kv_indptr, a->o, a->lse, a->sm_scale,
                               a->v_scale, a->logits_soft_cap, a->pos_encoding_mode, a->alibi_slopes,
                               a->rope_scale, a->rope_theta, a->sinks, a->enable_pdl, a->device, a->stream,
                               a->q_rows > 1 ? a->q_rows : 1, a->q_rows_noncausal);
}

int fi_batch_decode_step(const FiDecodePlanArgs* plan, const FiDecodeRunArgs* run) {
  if (plan == nullptr || run == nullptr) return fail(FI_ERR_INVALID_ARGUMENT, "null argument block");
  if (run->plan_info != plan->plan_info_out)
    return fail(FI_ERR_INVALID_ARGUMENT, "fi_batch_decode_step: run->plan_info must be plan->plan_info_out");
  if (int rc = fi_batch_decode_plan_args(plan)) return rc;
  if (plan->kv_indices_host != nullptr && run->kv_indices == nullptr && !plan->cuda_graph &&
      plan->plan_info_out[kPiOffIndices] < 0)
    return fail(FI_ERR_WORKSPACE, "fi_batch_decode_step: the page table does not fit the int workspace and no "
                                  "device copy was given");
  return fi_batch_decode_run_args(run);
}

size_t fi_single_decode_tmp_bytes(int num_qo_heads, int num_kv_heads, int head_dim, int device) {
  int sms = 148;
  if (device >= 0) device_sm_count(device, &sms);
  const size_t slots = (size_t)sms * 2 + num_kv_heads;
  const size_t g = num_qo_heads / (num_kv_heads > 0 ? num_kv_heads : 1);
  return slots * g * (head_dim + 1) * 4 + 64;
}

int fi_single_decode_run(const void* q, const void* k, const void* v, void* tmp, size_t tmp_bytes,
                         void* o, float* lse, int num_qo_heads, int num_kv_heads, int head_dim,
                         int kv_len, int q_dtype, int kv_dtype, int kv_layout, int window_left,
                         float sm_scale, float v_scale, float logits_soft_cap,
                         int pos_encoding_mode, const float* alibi_slopes, float rope_scale,
                         float rope_theta, int device, void* stream_) {
  if (int rc = check_pos_mode(pos_encoding_mode, alibi_slopes, rope_scale, rope_theta)) return rc;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (num_kv_heads <= 0 || num_qo_heads % num_kv_heads)
    return fail(FI_ERR_INVALID_ARGUMENT, "num_qo_heads must be a multiple of num_kv_heads");
  const int kvf = kv_format(q_dtype, kv_dtype);
  if (kvf < 0) return fail(FI_ERR_UNSUPPORTED, "unsupported q/kv dtype pair");
  // GQA groups of 9 .. 16 as one 16-column unit (consume_rows2) where the batch path does it too:
  // 16-bit KV, head_dim <= 128, TMA loader (K / V 16-byte aligned, at least one tile), no ROPE_LLAMA
  const int full_group = num_qo_heads / num_kv_heads;
  const bool wide = kvf == 0 && head_dim <= 128 && pos_encoding_mode != FI_POS_ROPE_LLAMA && full_group > kMaxGroup &&
                    tuning().loader != 2 && kv_len >= tile_tokens(kvf) &&
                    ((reinterpret_cast<uintptr_t>(k) | reinterpret_cast<uintptr_t>(v)) & 15) == 0;
  const int kv_div = group_split(full_group, wide ? 2 * kMaxGroup : max_group_for(head_dim));
  const int num_kv_heads_mem = num_kv_heads;  // heads k / v really have
  num_kv_heads *= kv_div;                     // virtual heads: one unit each (see DecodeParams)
  const int group = num_qo_heads / num_kv_heads;
  if (!(head_dim == 64 || head_dim == 128 || head_dim == 256 || head_dim == 512))
    return fail(FI_ERR_UNSUPPORTED, "unsupported head_dim %d", head_dim);
  if (kv_len < 0) return fail(FI_ERR_INVALID_ARGUMENT, "negative kv_len");
  const int es = dtype_size(kv_dtype);
  FI_CUDA_TRY(cudaSetDevice(device));
  int sms = 0;
  int rc = device_sm_count(device, &sms);
  if (rc != FI_OK) return rc;
  const KernelConfig cfg = choose_config(kvf, head_dim);

  const int tok_lo = window_left >= 0 ? (kv_len - 1 - window_left > 0 ? kv_len - 1 - window_left : 0) : 0;
  const int TT = tile_tokens(kvf);
  const int tiles = kv_len == 0 ? 0 : (kv_len + TT - 1) / TT - tok_lo / TT;
  const long long T = (long long)tiles * num_kv_heads;
  const int min_tiles = tuning().min_tiles_per_cta > 0 ? tuning().min_tiles_per_cta : 4;
  long long grid = T / min_tiles;
  if (grid < 1) grid = 1;
  if (grid > (long long)sms * cfg.ctas_per_sm) grid = (long long)sms * cfg.ctas_per_sm;

  const size_t slots = (size_t)grid + num_kv_heads;
  const size_t tmp_o_bytes = ((slots * group * head_dim * 4) + 15) & ~size_t(15);
  if (tmp_o_bytes + slots * group * 4 > tmp_bytes)
    return fail(FI_ERR_WORKSPACE, "single decode tmp buffer too small");

  DecodeParams p;
  memset(&p, 0, sizeof(p));
  p.q = q; p.q_stride_n = 0; p.q_stride_h = head_dim;
  p.o = o; p.lse = lse; p.k = k; p.v = v;
  p.kv_head_div = kv_div;
  p.indices_early = 1;
  if (kv_layout == FI_LAYOUT_NHD) { p.kv_stride_n = (int64_t)num_kv_heads_mem * head_dim; p.kv_stride_h = head_dim; }
  else { p.kv_stride_n = head_dim; p.kv_stride_h = (int64_t)kv_len * head_dim; }
  p.kv_stride_page = 0;
  p.page_size = kv_len > 0 ? kv_len : 1;
  p.page_shift = 31; p.page_mask = 0x7fffffff;
  p.rows_per_box = TT;
  p.tma_h_first = p.kv_stride_h < p.kv_stride_n;
  p.header = nullptr;
  p.imp_units = num_kv_heads; p.imp_tiles = tiles; p.imp_tok_lo = tok_lo; p.imp_tok_hi = kv_len;
  p.tmp_o = static_cast<float*>(tmp);
  p.tmp_lse = reinterpret_cast<float*>(static_cast<char*>(tmp) + tmp_o_bytes);
  p.num_qo_heads = num_qo_heads; p.group_size = group;
  p.num_stages = cfg.stages;
  p.sm_scale = sm_scale; p.sm_scale_log2 = sm_scale * 1.4426950408889634f;
  p.v_scale = v_scale; p.soft_cap = logits_soft_cap; p.enable_pdl = 0; p.l2_hints = default_l2_hints();
  p.alibi_slopes = pos_encoding_mode == FI_POS_ALIBI ? alibi_slopes : nullptr;
  p.rope = pos_encoding_mode == FI_POS_ROPE_LLAMA;
  p.rope_rcp_scale = 1.f / rope_scale; p.rope_rcp_theta = 1.f / rope_theta;
  if (group > kMaxGroup) { p.q_rows = 1; p.q_group = full_group; }  // wide units (see above)

  const bool tma = !(kvf != 0 && head_dim == 64) && head_dim != 512 &&
                   tma_eligible(k, v, es, p.kv_stride_n, p.kv_stride_h, 0, 0, true, kv_len, TT);
  CUtensorMap km, vm;
  memset(&km, 0, sizeof(km));
  memset(&vm, 0, sizeof(vm));
  if (tma) {
    const int64_t sp = (int64_t)kv_len * num_kv_heads_mem * head_dim;  // single "page"
    rc = make_kv_map(&km, k, kv_dtype, head_dim, kv_len, num_kv_heads_mem, 1, p.kv_stride_n, p.kv_stride_h, sp, TT);
    if (rc != FI_OK) return rc;
    rc = make_kv_map(&vm, v, kv_dtype, head_dim, kv_len, num_kv_heads_mem, 1, p.kv_stride_n, p.kv_stride_h, sp, TT);
    if (rc != FI_OK) return rc;
  }
  rc = launch_decode(p, km, vm, q_dtype, kvf, head_dim, (int)grid, tma, stream);
  if (rc != FI_OK) return rc;
  g_last_loader = tma ? 0 : 1;
  ++g_launch_count;
  if (grid > 1 && T > 0) {
    MergeParams mp;
    memset(&mp, 0, sizeof(mp));
    mp.header = nullptr;
    mp.imp_units = num_kv_heads; mp.imp_tiles = tiles; mp.imp_ctas = (int)grid;
    mp.tmp_o = p.tmp_o; mp.tmp_lse = p.tmp_lse; mp.o = o; mp.lse = lse;
    mp.num_qo_heads = num_qo_heads; mp.group_size = group; mp.head_dim = head_dim;
    mp.out_scale = v_scale; mp.out_dtype = q_dtype; mp.enable_pdl = 0;
    rc = launch_merge(mp, num_kv_heads, stream);
    if (rc != FI_OK) return rc;
    ++g_launch_count;
  }
  return FI_OK;
}

}  // extern "C"

// ---- allocation-free single decode (no reference counterpart; see fi_b200.h) -------------------
// The split-KV partials of fi_single_decode_run_static live in this file-scope array.  In host code
// `g_fi_static_tmp` is only a stand-in for the symbol, not a device address: the pointer the
// kernels get comes from cudaGetSymbolAddress, taken once per device with that device current.
static __device__ __align__(256) unsigned char g_fi_static_tmp[FI_SINGLE_DECODE_STATIC_TMP_BYTES];

namespace fib200 {
static int static_tmp_address(int device, void** out) {
  static std::mutex mu;
  static void* addr[64];
  if (device < 0 || device >= 64) return fail(FI_ERR_INVALID_ARGUMENT, "bad device %d", device);
  std::lock_guard<std::mutex> lk(mu);
  if (addr[device] == nullptr) {
    FI_CUDA_TRY(cudaSetDevice(device));
    void* p = nullptr;
    FI_CUDA_TRY(cudaGetSymbolAddress(&p, g_fi_static_tmp));
    if (p == nullptr) return fail(FI_ERR_CUDA, "cudaGetSymbolAddress returned a null scratch address");
    addr[device] = p;
  }
  *out = addr[device];
  return FI_OK;
}
}  // namespace fib200

extern "C" int fi_single_decode_run_static(const void* q, const void* k, const void* v, void* o,
                                           float* lse, int num_qo_heads, int num_kv_heads,
                                           int head_dim, int kv_len, int q_dtype, int kv_dtype,
                                           int kv_layout, int window_left, float sm_scale,
                                           float v_scale, float logits_soft_cap,
                                           int pos_encoding_mode, const float* alibi_slopes,
                                           float rope_scale, float rope_theta, int device,
                                           void* stream) {
  if (num_kv_heads <= 0 || num_qo_heads <= 0 || num_qo_heads % num_kv_heads)
    return fail(FI_ERR_INVALID_ARGUMENT, "num_qo_heads must be a multiple of num_kv_heads");
  const size_t need = fi_single_decode_tmp_bytes(num_qo_heads, num_kv_heads, head_dim, device);
  if (need > (size_t)FI_SINGLE_DECODE_STATIC_TMP_BYTES)
    return fail(FI_ERR_WORKSPACE,
                "fi_single_decode_run_static: %d/%d heads at head_dim %d need %zu bytes of scratch, the "
                "static array holds %zu (FI_SINGLE_DECODE_STATIC_TMP_BYTES); use fi_single_decode_run "
                "with a caller-owned tmp buffer",
                num_qo_heads, num_kv_heads, head_dim, need, (size_t)FI_SINGLE_DECODE_STATIC_TMP_BYTES);
  void* tmp = nullptr;
  if (int rc = static_tmp_address(device, &tmp)) return rc;
  // nothing in the scratch has to start at zero (every partial slot is written by the decode launch
  // before the merge launch reads it), so the previous call's contents are simply overwritten
  return fi_single_decode_run(q, k, v, tmp, (size_t)FI_SINGLE_DECODE_STATIC_TMP_BYTES, o, lse,
                              num_qo_heads, num_kv_heads, head_dim, kv_len, q_dtype, kv_dtype,
                              kv_layout, window_left, sm_scale, v_scale, logits_soft_cap,
                              pos_encoding_mode, alibi_slopes, rope_scale, rope_theta, device, stream);
}
