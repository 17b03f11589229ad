/*
 * fi_b200.h — C-ABI boundary of the B200-native paged-KV batch decode path.
 *
 * Every entry point replaces one tvm-ffi typed function of the reference
 * (happierpig/flashinfer-ai, FlashInfer 0.3.1).  Plain pointers and sizes only:
 * no torch / DLPack / tvm types cross this boundary.  All device memory is
 * caller-owned; the library never allocates device memory.
 *
 * Return value: 0 on success, non-zero FI_ERR_* otherwise; fi_last_error()
 * returns a thread-local human-readable message for the last failure.
 *
 * LSE values are base-2 (m + log2(d) with logits pre-multiplied by
 * sm_scale*log2(e)), as in the reference (include/flashinfer/attention/state.cuh:45).
 */
#ifndef FI_B200_H_
#define FI_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define FI_API __attribute__((visibility("default")))
#else
#define FI_API
#endif

/* ---- status codes ------------------------------------------------------- */
#define FI_OK 0
#define FI_ERR_INVALID_ARGUMENT 1 /* maps to Python ValueError            */
#define FI_ERR_WORKSPACE 2        /* workspace too small -> RuntimeError  */
#define FI_ERR_CUDA 3             /* CUDA runtime failure -> RuntimeError */
#define FI_ERR_UNSUPPORTED 4      /* shape / dtype not built -> RuntimeError */
#define FI_ERR_REPLAN 5           /* the plan's unit geometry does not fit this run() (see
                                     FI_PLAN_HEAD_PAIRS): plan again with flags = 0 and retry */

/* ---- dtype codes -------------------------------------------------------- */
#define FI_DTYPE_F16 0
#define FI_DTYPE_BF16 1
#define FI_DTYPE_E4M3 2
#define FI_DTYPE_E5M2 3
#define FI_DTYPE_F32 4 /* merge operators only */

/* ---- kv layout codes (reference: flashinfer/utils.py:35-38 TensorLayout) - */
/* position encoding applied inside the kernel: the reference's PosEncodingMode
 * (include/flashinfer/pos_enc.cuh:34-41, flashinfer/utils.py:35-38) */
#define FI_POS_NONE 0
#define FI_POS_ROPE_LLAMA 1
#define FI_POS_ALIBI 2

#define FI_LAYOUT_NHD 0
#define FI_LAYOUT_HND 1

/* number of int64 slots in the plan-info vector (reference DecodePlanInfo has
 * 10, include/flashinfer/attention/scheduler.cuh:366-422; ours carries the
 * persistent-grid work list offsets instead and is opaque to Python). */
#define FI_PLAN_INFO_LEN 32

FI_API const char* fi_last_error(void);
FI_API int fi_version(void);

/* Number of int64 plan-info slots (== FI_PLAN_INFO_LEN). */
FI_API int fi_plan_info_len(void);

/*
 * Process-wide tuning overrides used by benchmarks / tests (0 = automatic):
 *  stages             cap on the shared-memory ring depth
 *  ctas_per_sm        persistent CTAs per SM (1 or 2)
 *  loader             0 auto, 1 force per-page TMA boxes, 2 force cp.async gathers
 *  min_tiles_per_cta  smallest per-CTA share (64-token tiles) before CTAs are left idle
 * No reference counterpart (the reference fixes these at JIT time,
 * include/flashinfer/attention/decode.cuh:752-772).
 */
FI_API void fi_set_tuning(int stages, int ctas_per_sm, int loader, int min_tiles_per_cta);
/* loader used by the last decode launch: 0 = TMA, 1 = cp.async, -1 = none yet */
FI_API int fi_last_loader(void);
/* number of kernels launched by this library so far (decode + merge + merge ops) */
FI_API long long fi_launch_count(void);

/*
 * fi_batch_decode_plan — host-side scheduler.
 * Replaces: BatchDecodeWithPagedKVCachePlan (csrc/batch_decode.cu:39-79) ->
 *           DecodePlan (include/flashinfer/attention/scheduler.cuh:424-493).
 *
 * Cuts the flattened (request, kv-head, tile) stream into contiguous work items (by
 * default `num_ctas` equal ranges, one per persistent CTA), emits item heads, segments
 * and the split-KV merge metadata into `pinned_int_ws` (host) and issues an async H2D
 * copy into `int_ws` (device) on `stream`.
 *
 *  float_ws/float_ws_bytes   device scratch for split-KV partials (caller-owned)
 *  int_ws/pinned_int_ws      device / page-locked host mirrors, int_ws_bytes each
 *  indptr_host[bs+1]         page indptr (host)
 *  last_page_len_host[bs]    host; may be NULL iff kv_len_host != NULL
 *  kv_len_host[bs]           optional explicit token lengths (host) or NULL
 *  window_left               <0: none; else only the last window_left+1 tokens
 *  cuda_graph                1: the grid and every offset run() passes to the kernels are
 *                            fixed by (bs, heads, num_ctas) only, so re-plans are graph-safe
 *  fixed_split_pages         >0: batch-invariant split (reference `fixed_split_size`,
 *                            decode.py:880-885): every request is cut at multiples of
 *                            this many pages from its start, chunks are never subdivided
 *                            or fused, requests with more than one chunk are merged in
 *                            chunk order; <=0: off
 *  disable_split_kv          !=0: a (request, kv head) unit is never split (reference
 *                            decode.py:886-887)
 *  num_ctas_override         >0: use this many CTAs along x instead of querying
 *                            the device (used by CPU-only tests; then `device`
 *                            may be -1 and no CUDA call is made: the H2D copy
 *                            becomes a memcpy into int_ws)
 *  kv_indices_host[num_indices]  optional: the page table on the host (NULL: unknown to the
 *                            planner).  When given, (a) the page ids of every work item's first tiles
 *                            are written into its head, so the kernel issues its first TMA boxes
 *                            without a page-table lookup, (b) outside CUDA-graph mode the table is
 *                            appended to the work list and travels in the same H2D copy: run() may
 *                            then be called with kv_indices = NULL, (c) the kernel may read the table
 *                            ahead of its programmatic-dependent-launch wait (it is only ever written
 *                            by host copies).  The reference copies the table to the device in its
 *                            Python wrapper (flashinfer/decode.py:1009-1012) and never shows it to
 *                            the planner.
 *  plan_info_out[FI_PLAN_INFO_LEN]
 */
FI_API int fi_batch_decode_plan(void* float_ws, size_t float_ws_bytes, void* int_ws,
                         void* pinned_int_ws, size_t int_ws_bytes,
                         const int32_t* indptr_host,
                         const int32_t* last_page_len_host,
                         const int32_t* kv_len_host, int batch_size,
                         int num_qo_heads, int num_kv_heads, int head_dim,
                         int page_size, int q_dtype, int kv_dtype,
                         int window_left, int cuda_graph, int fixed_split_pages,
                         int disable_split_kv, int device, int num_ctas_override,
                         void* stream, const int32_t* kv_indices_host, int64_t num_indices,
                         int64_t* plan_info_out);

/*
 * fi_batch_decode_run — launches the decode kernel; split requests are merged
 * inside it (FI_B200_FUSED_MERGE=0 at plan time: by a second merge launch).
 * Replaces: BatchDecodeWithPagedKVCacheRun (csrc/batch_decode.cu:81-191) ->
 *           BatchDecodeWithPagedKVCacheDispatched
 *           (include/flashinfer/attention/decode.cuh:739-825).
 *
 *  q            [bs, Hq, D], element strides q_stride_n / q_stride_h, unit on D
 *  k_cache/v_cache  page pools of num_pages pages sharing element strides
 *                   (page, n, h), unit on D (reference constraint
 *                   csrc/batch_decode.cu:123-128)
 *  kv_indices   device int32 page table; NULL = the copy plan() made from kv_indices_host, or,
 *               when plan() was not given one, the identity (page id == position)
 *  kv_indptr    device int32 [bs+1], or NULL: the kernels take every request's page
 *               base from the planned work list (plan() read indptr on the host), the
 *               argument is kept for signature parity with the reference run()
 *  o            [bs, Hq, D] contiguous, q dtype
 *  lse          [bs, Hq] f32 base-2 or NULL
 *  sm_scale     already multiplied by q_scale*k_scale (decode.py:1250-1256)
 *  v_scale      multiplies the output (fused; reference: decode.py:1367-1372)
 *  logits_soft_cap  <=0: disabled
 *  pos_encoding_mode  FI_POS_NONE / FI_POS_ROPE_LLAMA / FI_POS_ALIBI
 *  alibi_slopes     device f32 [Hq], read in FI_POS_ALIBI mode (variants.cuh:68-71 with
 *                   qo_idx = 0 as passed by decode.cuh:95-96):
 *                   logit = q.k * sm_scale + alibi_slopes[head] * kv_position
 *  rope_scale, rope_theta  FI_POS_ROPE_LLAMA: q is rotated at position kv_len - 1 and every K
 *                   row at its own position, frequencies (1 / rope_scale) * rope_theta^(-2i/D)
 *                   (decode.cuh:445-466, pos_enc.cuh:79-100; Python defaults 1.0 and 1e4,
 *                   decode.py:1257-1260).  16-bit KV: K rotated in registers inside the MMA kernel;
 *                   fp8 KV: a CUDA-core variant kernel.
 *  sinks            device f32 [Hq] or NULL (reference run(..., sinks), decode.py:1176): one extra
 *                   logit per query head in the softmax denominator only,
 *                   p_i = exp(s_i) / (sum_j exp(s_j) + exp(sinks[head])); lse includes it
 */
FI_API int fi_batch_decode_run(void* float_ws, void* int_ws, const int64_t* plan_info,
                        const void* q, int64_t q_stride_n, int64_t q_stride_h,
                        const void* k_cache, const void* v_cache,
                        int64_t num_pages, int64_t kv_stride_page, int64_t kv_stride_n,
                        int64_t kv_stride_h, const int32_t* kv_indices,
                        const int32_t* kv_indptr, void* o, float* lse,
                        float sm_scale, float v_scale, float logits_soft_cap,
                        int pos_encoding_mode, const float* alibi_slopes, float rope_scale,
                        float rope_theta, const float* sinks, int enable_pdl, int device,
                        void* stream);

/*
 * The same two calls with their arguments in a caller-owned block: a serving loop fills the block
 * once per (wrapper, cache) and rewrites only the pointers that change per step / per layer, so the
 * per-step cost of the binding (ctypes here, cgo / JNI elsewhere) is one pointer argument instead
 * of 26.  Field for field the arguments of fi_batch_decode_plan / fi_batch_decode_run above; no
 * reference counterpart (tvm-ffi packs its arguments per call, csrc/batch_decode_jit_binding.cu:23-40).
 */
/*
 * FI_PLAN_HEAD_PAIRS: the caller's cache is NHD (the rows of consecutive KV heads of a token are
 * adjacent) and run() will not ask for ROPE_LLAMA.  With fp8 KV at head_dim 128, an even number of
 * KV heads, a GQA group <= 8 and a power-of-two page size >= 8 the planner then makes (request, PAIR
 * of KV heads) units for the tcgen05 kernel, which streams both heads' 128-byte rows of a token as
 * one 256-byte run (C3 on one B200: 0.92 -> 0.9x of the measured copy bandwidth); every other shape
 * ignores the bit.  A head-pair plan run on a cache the TMA loader cannot address (strides that are
 * not multiples of 16 bytes), or with ROPE_LLAMA, returns FI_ERR_REPLAN.  No reference counterpart:
 * the reference's planner knows nothing of the cache layout (scheduler.cuh:411-481).
 */
#define FI_PLAN_HEAD_PAIRS 1
/*
 * FI_PLAN_WIDE_UNITS: units of up to 16 query-head columns instead of 8 (16-bit KV at head_dim <= 128 or fp8 KV at 128,
 * power-of-two pages >= 8; the caller promises that run() will not ask for ROPE_LLAMA).  The 16-bit
 * kernel then carries two 8-column MMA tiles per KV pass: GQA groups of 9 .. 16 (Llama-3-405B: 16)
 * stream their KV head once instead of once per 8 query heads, and plans that run with
 * FiDecodeRunArgs.q_rows > 1 (multi-token decode, cascade levels) fit 16 (row, head) pairs in a
 * unit.  Such a plan run with ROPE_LLAMA or through the gather loader returns FI_ERR_REPLAN.
 * No reference counterpart.
 */
#define FI_PLAN_WIDE_UNITS 2
typedef struct FiDecodePlanArgs {
  void* float_ws; size_t float_ws_bytes;
  void* int_ws; void* pinned_int_ws; size_t int_ws_bytes;
  const int32_t* indptr_host; const int32_t* last_page_len_host; const int32_t* kv_len_host;
  int32_t batch_size, num_qo_heads, num_kv_heads, head_dim, page_size, q_dtype, kv_dtype;
  int32_t window_left, cuda_graph, fixed_split_pages, disable_split_kv, device, num_ctas_override;
  int32_t flags;   /* FI_PLAN_* bits; the positional fi_batch_decode_plan plans with 0 */
  void* stream;
  const int32_t* kv_indices_host; int64_t num_indices;
  int64_t* plan_info_out;
} FiDecodePlanArgs;
FI_API int fi_batch_decode_plan_args(const FiDecodePlanArgs* a);

typedef struct FiDecodeRunArgs {
  void* float_ws; void* int_ws; const int64_t* plan_info;
  const void* q; int64_t q_stride_n, q_stride_h;
  const void* k_cache; const void* v_cache;
  int64_t num_pages, kv_stride_page, kv_stride_n, kv_stride_h;
  const int32_t* kv_indices; const int32_t* kv_indptr;
  void* o; float* lse;
  float sm_scale, v_scale, logits_soft_cap; int32_t pos_encoding_mode;
  const float* alibi_slopes; float rope_scale, rope_theta;
  const float* sinks;
  int32_t enable_pdl, device;
  void* stream;
  /* Multi-token decode (argument-block form only; <= 1: one query row per request).  q / o hold
   * q_rows consecutive rows per request ([batch * q_rows, Hq, D], q_stride_n = the row stride; lse
   * [batch * q_rows, Hq]); row r of a request attends all but the last q_rows - 1 - r tokens of its
   * cache (causal among the new tokens, whose K / V the cache already holds).  The plan must have
   * been made with num_qo_heads = q_rows * Hq on the request's own page table: the rows then share
   * one KV stream as extra columns of the GQA group instead of re-reading it once per row.
   * 16-bit KV, NONE / ALIBI, no sliding window; anything else returns FI_ERR_UNSUPPORTED.  Reference:
   * q_len_per_req of BatchDecodeWithPagedKVCacheWrapper.run (flashinfer/decode.py:1177,1278).
   * q_rows_noncausal = 1: every row sees the whole cache (the shared levels of cascade attention,
   * flashinfer/cascade.py:498-543: the rows of a group attend the group's pages without a mask). */
  int32_t q_rows, q_rows_noncausal;
} FiDecodeRunArgs;
FI_API int fi_batch_decode_run_args(const FiDecodeRunArgs* a);
/*
 * fi_batch_decode_step — fi_batch_decode_plan_args(plan) followed by fi_batch_decode_run_args(run) in ONE call:
 * a decode step of a serving loop (new page table, same problem shape) costs the binding a single
 * foreign call.  run->plan_info must be plan->plan_info_out.  When the planner could not place its copy
 * of kv_indices_host behind the work list (int workspace too small) and run->kv_indices is NULL the
 * call fails with FI_ERR_WORKSPACE before anything is launched (NULL would mean "identity table").
 * No reference counterpart: its wrapper calls plan and run through separate tvm-ffi functions
 * (flashinfer/decode.py:1086-1102, 1286-1330).
 */
FI_API int fi_batch_decode_step(const FiDecodePlanArgs* plan, const FiDecodeRunArgs* run);

/*
 * fi_single_decode_run — one request, contiguous K/V.
 * Replaces: single_decode_with_kv_cache FFI (csrc/single_decode.cu:33-104) ->
 *           SingleDecodeWithKVCacheDispatched (decode.cuh:658-737).
 *  k, v   NHD [L, Hkv, D] or HND [Hkv, L, D], contiguous
 *  tmp    device scratch (>= fi_single_decode_tmp_bytes) for split-KV partials
 *         and the work list; int_host_scratch is not needed: the plan for a
 *         single request is closed-form and computed on the device.
 */
FI_API int fi_single_decode_run(const void* q, const void* k, const void* v, void* tmp,
                         size_t tmp_bytes, void* o, float* lse, int num_qo_heads,
                         int num_kv_heads, int head_dim, int kv_len,
                         int q_dtype, int kv_dtype, int kv_layout,
                         int window_left, float sm_scale, float v_scale,
                         float logits_soft_cap, int pos_encoding_mode,
                         const float* alibi_slopes, float rope_scale, float rope_theta,
                         int device, void* stream);
FI_API size_t fi_single_decode_tmp_bytes(int num_qo_heads, int num_kv_heads, int head_dim,
                                  int device);

/*
 * fi_single_decode_run_static — fi_single_decode_run without the caller's `tmp`.
 * No counterpart in the reference: added for callers
 * that cannot allocate device memory (CUDA-graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  The split-KV partials go to a file-scope `static __device__` array of
 * FI_SINGLE_DECODE_STATIC_TMP_BYTES inside the library (every GPU that loads the library holds
 * it until the process exits); its address is taken with cudaGetSymbolAddress once per device.
 * The call allocates nothing: no cudaMalloc / cudaFree / allocating library routine.
 *  - shapes whose scratch need (fi_single_decode_tmp_bytes) exceeds the array are rejected with
 *    FI_ERR_WORKSPACE and a message naming both sizes;
 *  - every call on a device shares the one array, which keeps the previous call's contents
 *    (nothing in it has to start at zero: each launch writes every partial slot before its merge
 *    reads it); calls that overlap on the GPU (two streams, or two host threads) race on it —
 *    serialise them on one stream.
 * All other arguments as fi_single_decode_run.
 */
#define FI_SINGLE_DECODE_STATIC_TMP_BYTES (4u << 20)
FI_API int fi_single_decode_run_static(const void* q, const void* k, const void* v, void* o,
                                float* lse, int num_qo_heads, int num_kv_heads, int head_dim,
                                int kv_len, int q_dtype, int kv_dtype, int kv_layout,
                                int window_left, float sm_scale, float v_scale,
                                float logits_soft_cap, int pos_encoding_mode,
                                const float* alibi_slopes, float rope_scale, float rope_theta,
                                int device, void* stream);

/*
 * State-merge operators.  Replaces csrc/cascade.cu:23 / :59 / :102 ->
 * MergeState / MergeStateInPlace / MergeStates
 * (include/flashinfer/attention/cascade.cuh:573 / :605 / :638).
 *  v_*  [seq, H, D] contiguous, dtype in {F16, BF16} (or F32 accepted too)
 *  s_*  [seq, H] f32 base-2 LSE
 */
FI_API int fi_merge_state(const void* v_a, const float* s_a, const void* v_b,
                   const float* s_b, void* v_merged, float* s_merged,
                   int seq_len, int num_heads, int head_dim, int dtype,
                   int device, void* stream);
FI_API int fi_merge_state_in_place(void* v, float* s, const void* v_other,
                            const float* s_other, const uint8_t* mask_or_null,
                            int seq_len, int num_heads, int head_dim, int dtype,
                            int device, void* stream);
/* v [seq, n, H, D], s [seq, n, H] -> v_merged [seq, H, D], s_merged [seq, H] */
FI_API int fi_merge_states(const void* v, const float* s, void* v_merged,
                    float* s_merged, int seq_len, int num_index_sets,
                    int num_heads, int head_dim, int dtype, int device,
                    void* stream);
/*
 * Variable-length merge used after split-KV decode.
 * Replaces VariableLengthMergeStates (cascade.cuh:686-736).
 *  v [total, H, D] (dtype_in), s [total, H]; indptr device int32 [seq+1];
 *  rows with indptr[i+1]-indptr[i]==0 are left untouched if skip_empty!=0,
 *  else written as zeros / lse=-5e4 (reference cascade.cuh:397-405).
 */
FI_API int fi_variable_length_merge_states(const void* v, const float* s,
                                    const int32_t* indptr, void* v_merged,
                                    float* s_merged, int seq_len, int num_heads,
                                    int head_dim, int dtype_in, int dtype_out,
                                    int skip_empty, float out_scale, int device,
                                    void* stream);

/* ---- page table maintenance: the step before decode in every iteration -------------------- */
/*
 * fi_get_batch_indices_positions — ragged append indptr -> per-token (batch index, position).
 * Replaces the Triton get_batch_indices_positions_kernel launched by
 * flashinfer.get_batch_indices_positions (flashinfer/page.py:169-221,
 * flashinfer/triton/page.py:21-42):
 *   batch_indices[j] = i,  positions[j] = j + seq_lens[i] - append_indptr[i+1]
 *   for append_indptr[i] <= j < append_indptr[i+1].
 * All pointers device int32; batch_indices / positions have nnz entries.
 */
FI_API int fi_get_batch_indices_positions(const int32_t* append_indptr, const int32_t* seq_lens,
                                   int32_t* batch_indices, int32_t* positions, int batch_size,
                                   int nnz, int device, void* stream);
/*
 * fi_append_paged_kv_cache — scatter nnz new (token, head) rows of K and V into the paged cache.
 * Replaces append_paged_kv_cache (csrc/page.cu:24-110) -> AppendPagedKVCache /
 * AppendPagedKVCacheKernel (include/flashinfer/page.cuh:254-287,370-420).
 *  append_key/value  [nnz, Hkv, D], element strides (n, h), unit stride on D (may be views)
 *  k_cache/v_cache   page pools with element strides (page, n, h) as in fi_batch_decode_run
 *  row j goes to page kv_indices[kv_indptr[b] + pos / page_size], entry pos % page_size,
 *  b = batch_indices[j], pos = positions[j] (the pages must already be allocated)
 *  dtype: element type of BOTH the appended rows and the cache (any FI_DTYPE_*: rows are copied
 *  as bytes, the reference requires the two dtypes to be equal as well, csrc/page.cu:45-52)
 */
FI_API int fi_append_paged_kv_cache(const void* append_key, const void* append_value,
                             int64_t ak_stride_n, int64_t ak_stride_h, int64_t av_stride_n,
                             int64_t av_stride_h, const int32_t* batch_indices,
                             const int32_t* positions, int64_t nnz, void* k_cache, void* v_cache,
                             int64_t kv_stride_page, int64_t kv_stride_n, int64_t kv_stride_h,
                             const int32_t* kv_indices, const int32_t* kv_indptr, int num_kv_heads,
                             int head_dim, int page_size, int dtype, int device, void* stream);

/* ---- multi-GPU: gather of the per-rank output heads over NVLink -------------------------- */
/*
 * fi_peer_gather — store this rank's rows into the gather buffer of EVERY rank and wait until every
 * rank's rows have arrived in ours.  The reference's decode path has no multi-GPU step (frameworks on
 * top of it all-gather O with NCCL); this replaces that all-gather + transpose for KV-head sharding
 * (flashinfer_b200/sharding.py PeerGather; protocol studied in flashinfer/comm/cuda_ipc.py and
 * flashinfer/comm/vllm_ar.py:105-155).
 *  src              this rank's rows [rows, row_bytes] (e.g. o_local [bs, Hq/N, D]), device memory
 *  peer_dst[world]  HOST array: rank p's destination buffer as mapped in this process (peer memory:
 *                   CUDA IPC / symmetric memory); row r of this rank lands at
 *                   peer_dst[p] + dst_offset_bytes + r * dst_row_stride_bytes
 *  mc_dst           NVLS multicast address of the same buffer or NULL: when set (and everything is
 *                   16-byte aligned) one multimem.st per chunk replaces the `world` unicast stores
 *  peer_ctl[world]  HOST array: rank p's 256-byte control block as mapped here (zero at creation:
 *                   uint32 flag[16], CTA counter, status word)
 *  epoch            must grow by one per call on every rank (same value on all ranks for one gather)
 * The kernel is chained to the preceding launch of `stream` with programmatic dependent launch when
 * enable_pdl != 0.  A peer that never arrives makes the wait time out after 2 s (status word != 0).
 */
FI_API int fi_peer_gather(const void* src, void* const* peer_dst, void* mc_dst,
                          void* const* peer_ctl, int world, int rank, int64_t rows,
                          int64_t row_bytes, int64_t dst_row_stride_bytes, int64_t dst_offset_bytes,
                          uint32_t epoch, int enable_pdl, int device, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* FI_B200_H_ */
