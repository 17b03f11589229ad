// oracle/ref_shim.cu — TEST INFRASTRUCTURE, NOT PRODUCT.
//
// Plain extern "C" entry points over the REFERENCE's own header-only implementation of the
// hot path, compiled where the headers lie (-I /root/reference/include, see oracle/Makefile)
// into oracle/_ref/liboracle_ref.so.  No reference source is copied into this repo: this file
// only instantiates the reference templates
//   DecodePlan                             include/flashinfer/attention/scheduler.cuh:424-493
//   BatchDecodeWithPagedKVCacheDispatched  include/flashinfer/attention/decode.cuh:739-825
//   SingleDecodeWithKVCacheDispatched      include/flashinfer/attention/decode.cuh:658-737
//   MergeState / MergeStateInPlace / MergeStates / VariableLengthMergeStates
//                                          include/flashinfer/attention/cascade.cuh:573-736
// with parameter structs carrying the fields those templates read (the reference renders the
// same structs from csrc/batch_decode_customize_config.jinja and
// csrc/single_decode_customize_config.jinja at JIT time).
// Only tests/ and bench.py's comparator leg load the resulting library.
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_fp8.h>

#include <flashinfer/attention/cascade.cuh>
#include <flashinfer/attention/decode.cuh>
#include <flashinfer/attention/scheduler.cuh>
#include <flashinfer/attention/variants.cuh>
#include <flashinfer/page.cuh>
#include <flashinfer/pos_enc.cuh>

#include <cstdint>
#include <cstdio>
#include <string>

using namespace flashinfer;

namespace refshim {

thread_local std::string g_err;

template <typename Q, typename KV>
struct BatchParams {
  using DTypeQ = Q;
  using DTypeKV = KV;
  using DTypeO = Q;
  using IdType = int32_t;
  DTypeQ* q;
  paged_kv_t<DTypeKV, IdType> paged_kv;
  DTypeO* o;
  float* lse;
  float* maybe_alibi_slopes;
  float logits_soft_cap, sm_scale, rope_rcp_scale, rope_rcp_theta;
  uint32_t padded_batch_size, num_qo_heads;
  IdType q_stride_n, q_stride_h;
  int32_t window_left;
  bool enable_pdl;
  IdType *request_indices, *kv_tile_indices, *o_indptr, *kv_chunk_size_ptr;
  bool* block_valid_mask;
  bool partition_kv;
  __host__ __device__ __forceinline__ int32_t get_qo_len(int32_t) const { return 1; }
  __host__ __device__ __forceinline__ int32_t get_kv_len(int32_t b) const {
    return paged_kv.get_length(b);
  }
};

template <typename Q, typename KV>
struct SingleParams {
  using DTypeQ = Q;
  using DTypeKV = KV;
  using DTypeO = Q;
  using IdType = int32_t;
  DTypeQ* q;
  DTypeKV *k, *v;
  DTypeO* o;
  float* lse;
  float* maybe_alibi_slopes;
  float logits_soft_cap, sm_scale, rope_rcp_scale, rope_rcp_theta;
  uint32_t kv_len, num_qo_heads, num_kv_heads, q_stride_n, q_stride_h, kv_stride_n, kv_stride_h;
  int32_t window_left;
  uint32_t kv_chunk_size;
  __host__ __device__ __forceinline__ uint32_t get_qo_len(uint32_t) const { return 1; }
  __host__ __device__ __forceinline__ uint32_t get_kv_len(uint32_t) const { return kv_len; }
};

using Variant = DefaultAttention</*custom_mask=*/false, /*swa=*/false, /*softcap=*/false, /*alibi=*/false>;

struct BatchArgs {
  void *float_ws, *int_ws, *pinned_ws;
  size_t float_bytes, int_bytes;
  int32_t* indptr_h;
  int bs, hq, hkv, page_size, layout, cuda_graph;
  void *q, *k, *v, *o;
  float* lse;
  int64_t q_stride_n, q_stride_h, kv_strides[3];
  int32_t *indices, *indptr, *last_page_len;
  float sm_scale;
  cudaStream_t stream;
  int64_t* info_out;  // 10 x int64 DecodePlanInfo + [10] = launches (1 or 2)
  int mode;           // 0 plan + run, 1 plan only (kept for later runs), 2 run with the kept plan
  float rope_rcp_scale = 1.f, rope_rcp_theta = 1e-4f;  // read by the ROPE_LLAMA instantiation only
  // read by the logits-transform instantiations only (ref_batch_decode_variant)
  int32_t window_left = -1;
  float logits_soft_cap = 0.f;
  float* alibi_slopes = nullptr;
};

thread_local DecodePlanInfo g_plan;  // the reference keeps this in its Python wrapper between run()s
// Programmatic dependent launch: the reference's wrapper defaults enable_pdl to
// device_support_pdl(q.device), i.e. ON for sm_90+ (flashinfer/decode.py:1226-1227); the launch
// path is decode.cuh:775-821.  ref_set_pdl(0) switches it off for A/B runs.
int g_enable_pdl = 1;

template <typename Q, typename KV, uint32_t HD, PosEncodingMode POS = PosEncodingMode::kNone,
          typename VAR = Variant>
int batch_impl(const BatchArgs& a) {
  using P = BatchParams<Q, KV>;
  DecodePlanInfo& info = g_plan;
  cudaError_t st = cudaSuccess;
  bool ok = a.mode == 2;
  if (a.mode != 2) DISPATCH_GQA_GROUP_SIZE(a.hq / a.hkv, GROUP_SIZE, {
    auto est = BatchDecodeWithPagedKVCacheWorkEstimationDispatched<GROUP_SIZE, HD, POS, VAR, P>;
    st = DecodePlan<HD, POS, VAR, P>(a.float_ws, a.float_bytes, a.int_ws,
                                                            a.pinned_ws, a.int_bytes, info, a.indptr_h,
                                                            a.bs, a.hq, a.page_size, a.cuda_graph != 0,
                                                            a.stream, est);
    ok = true;
  });
  if (!ok) { g_err = "unsupported group size"; return 4; }
  if (st != cudaSuccess) { g_err = cudaGetErrorString(st); return 3; }
  if (a.mode == 1) return 0;
  paged_kv_t<KV, int32_t> kv(a.hkv, a.page_size, HD, a.bs, static_cast<QKVLayout>(a.layout),
                             static_cast<KV*>(a.k), static_cast<KV*>(a.v), a.kv_strides, a.indices,
                             a.indptr, a.last_page_len);
  P p;
  p.q = static_cast<Q*>(a.q);
  p.paged_kv = kv;
  p.o = static_cast<Q*>(a.o);
  p.lse = a.lse;
  p.maybe_alibi_slopes = a.alibi_slopes;
  p.logits_soft_cap = a.logits_soft_cap;
  p.sm_scale = a.sm_scale;
  p.rope_rcp_scale = a.rope_rcp_scale;
  p.rope_rcp_theta = a.rope_rcp_theta;
  p.padded_batch_size = info.padded_batch_size;
  p.num_qo_heads = a.hq;
  p.q_stride_n = a.q_stride_n;
  p.q_stride_h = a.q_stride_h;
  p.window_left = a.window_left;
  p.enable_pdl = g_enable_pdl != 0;
  char* ib = static_cast<char*>(a.int_ws);
  p.request_indices = reinterpret_cast<int32_t*>(ib + info.request_indices_offset);
  p.kv_tile_indices = reinterpret_cast<int32_t*>(ib + info.kv_tile_indices_offset);
  p.o_indptr = reinterpret_cast<int32_t*>(ib + info.o_indptr_offset);
  p.kv_chunk_size_ptr = reinterpret_cast<int32_t*>(ib + info.kv_chunk_size_ptr_offset);
  p.block_valid_mask = nullptr;
  p.partition_kv = false;
  Q* tmp_v = nullptr;
  float* tmp_s = nullptr;
  if (info.split_kv) {
    tmp_v = reinterpret_cast<Q*>(static_cast<char*>(a.float_ws) + info.v_offset);
    tmp_s = reinterpret_cast<float*>(static_cast<char*>(a.float_ws) + info.s_offset);
    if (info.enable_cuda_graph)
      p.block_valid_mask = reinterpret_cast<bool*>(ib + info.block_valid_mask_offset);
  }
  st = BatchDecodeWithPagedKVCacheDispatched<HD, POS, VAR>(p, tmp_v, tmp_s, g_enable_pdl != 0, a.stream);
  if (st != cudaSuccess) { g_err = cudaGetErrorString(st); return 3; }
  if (a.info_out) {
    auto v = info.ToVector();
    for (int i = 0; i < 10; ++i) a.info_out[i] = v[i];
    a.info_out[10] = info.split_kv ? 2 : 1;
  }
  return 0;
}

template <typename Q, typename KV, uint32_t HD>
int single_impl(void* q, void* k, void* v, void* tmp, void* o, float* lse, int hq, int hkv, int kv_len,
                int layout, float sm_scale, cudaStream_t stream) {
  SingleParams<Q, KV> p;
  p.q = static_cast<Q*>(q);
  p.k = static_cast<KV*>(k);
  p.v = static_cast<KV*>(v);
  p.o = static_cast<Q*>(o);
  p.lse = lse;
  p.maybe_alibi_slopes = nullptr;
  p.logits_soft_cap = 0.f;
  p.sm_scale = sm_scale;
  p.rope_rcp_scale = 1.f;
  p.rope_rcp_theta = 1e-4f;
  p.kv_len = kv_len;
  p.num_qo_heads = hq;
  p.num_kv_heads = hkv;
  p.q_stride_n = hq * HD;
  p.q_stride_h = HD;
  p.kv_stride_n = layout == 0 ? hkv * HD : HD;
  p.kv_stride_h = layout == 0 ? HD : kv_len * HD;
  p.window_left = -1;
  p.kv_chunk_size = 0;
  cudaError_t st = SingleDecodeWithKVCacheDispatched<HD, PosEncodingMode::kNone, Variant>(
      p, static_cast<Q*>(tmp), stream);
  if (st != cudaSuccess) { g_err = cudaGetErrorString(st); return 3; }
  return 0;
}

}  // namespace refshim
using namespace refshim;

// dtype codes as in include/fi_b200.h: 0 f16, 1 bf16, 2 e4m3, 3 e5m2 (f16 queries only)
#define REF_DISPATCH(qd, kvd, ...)                                                                  \
  if (qd == 0 && kvd == 0) { using Q = half; using KV = half; __VA_ARGS__; }                        \
  else if (qd == 1 && kvd == 1) { using Q = nv_bfloat16; using KV = nv_bfloat16; __VA_ARGS__; }     \
  else if (qd == 0 && kvd == 2) { using Q = half; using KV = __nv_fp8_e4m3; __VA_ARGS__; }          \
  else if (qd == 1 && kvd == 2) { using Q = nv_bfloat16; using KV = __nv_fp8_e4m3; __VA_ARGS__; }   \
  else if (qd == 0 && kvd == 3) { using Q = half; using KV = __nv_fp8_e5m2; __VA_ARGS__; }          \
  else { g_err = "unsupported dtype pair"; return 4; }

extern "C" {

__attribute__((visibility("default"))) const char* ref_last_error() { return g_err.c_str(); }
__attribute__((visibility("default"))) void ref_set_pdl(int on) { g_enable_pdl = on; }
__attribute__((visibility("default"))) int ref_get_pdl() { return g_enable_pdl; }

// plan (reference DecodePlan) + run (reference kernel [+ reference merge]); head_dim 128 only
__attribute__((visibility("default"))) int ref_batch_decode(
    void* float_ws, size_t float_bytes, void* int_ws, void* pinned_ws, size_t int_bytes,
    int32_t* indptr_h, int bs, int hq, int hkv, int head_dim, int page_size, int q_dtype,
    int kv_dtype, int layout, int cuda_graph, void* q, int64_t q_stride_n, int64_t q_stride_h,
    void* k, void* v, int64_t s0, int64_t s1, int64_t s2, int32_t* indices, int32_t* indptr,
    int32_t* last_page_len, void* o, float* lse, float sm_scale, void* stream, int64_t* info_out) {
  BatchArgs a{float_ws, int_ws, pinned_ws, float_bytes, int_bytes, indptr_h, bs, hq, hkv, page_size,
              layout, cuda_graph, q, k, v, o, lse, q_stride_n, q_stride_h, {s0, s1, s2}, indices,
              indptr, last_page_len, sm_scale, static_cast<cudaStream_t>(stream), info_out, 0};
  try {
    if (head_dim == 64 || head_dim == 256) {  // 16-bit KV only (keeps the build short)
      if (q_dtype == 0 && kv_dtype == 0)
        return head_dim == 64 ? batch_impl<half, half, 64>(a) : batch_impl<half, half, 256>(a);
      if (q_dtype == 1 && kv_dtype == 1)
        return head_dim == 64 ? batch_impl<nv_bfloat16, nv_bfloat16, 64>(a)
                              : batch_impl<nv_bfloat16, nv_bfloat16, 256>(a);
      g_err = "ref shim: head_dim 64 / 256 are built for 16-bit KV";
      return 4;
    }
    if (head_dim != 128) { g_err = "ref shim is built for head_dim 64 / 128 / 256"; return 4; }
    REF_DISPATCH(q_dtype, kv_dtype, return (batch_impl<Q, KV, 128>(a)));
  } catch (const std::exception& e) {
    g_err = e.what();
    return 2;
  }
}

// same, with plan and run separable as in the reference's Python wrapper (plan once, run per layer):
// mode 1 = plan only, 2 = run with the plan of the last mode-1 call on this thread, 0 = both
__attribute__((visibility("default"))) int ref_batch_decode_mode(
    void* float_ws, size_t float_bytes, void* int_ws, void* pinned_ws, size_t int_bytes,
    int32_t* indptr_h, int bs, int hq, int hkv, int head_dim, int page_size, int q_dtype,
    int kv_dtype, int layout, int cuda_graph, void* q, int64_t q_stride_n, int64_t q_stride_h,
    void* k, void* v, int64_t s0, int64_t s1, int64_t s2, int32_t* indices, int32_t* indptr,
    int32_t* last_page_len, void* o, float* lse, float sm_scale, void* stream, int64_t* info_out,
    int mode) {
  if (head_dim != 128) { g_err = "ref shim is built for head_dim 128"; return 4; }
  BatchArgs a{float_ws, int_ws, pinned_ws, float_bytes, int_bytes, indptr_h, bs, hq, hkv, page_size,
              layout, cuda_graph, q, k, v, o, lse, q_stride_n, q_stride_h, {s0, s1, s2}, indices,
              indptr, last_page_len, sm_scale, static_cast<cudaStream_t>(stream), info_out, mode};
  try {
    REF_DISPATCH(q_dtype, kv_dtype, return (batch_impl<Q, KV, 128>(a)));
  } catch (const std::exception& e) {
    g_err = e.what();
    return 2;
  }
}

// pos_encoding_mode = ROPE_LLAMA (reference decode.cuh:445-466: q rotated at kv_len - 1, every K
// row at its position), 16-bit KV only; same plan / run modes as ref_batch_decode_mode
__attribute__((visibility("default"))) int ref_batch_decode_rope(
    void* float_ws, size_t float_bytes, void* int_ws, void* pinned_ws, size_t int_bytes,
    int32_t* indptr_h, int bs, int hq, int hkv, int head_dim, int page_size, int q_dtype,
    int kv_dtype, int layout, int cuda_graph, void* q, int64_t q_stride_n, int64_t q_stride_h,
    void* k, void* v, int64_t s0, int64_t s1, int64_t s2, int32_t* indices, int32_t* indptr,
    int32_t* last_page_len, void* o, float* lse, float sm_scale, float rope_scale, float rope_theta,
    void* stream, int64_t* info_out, int mode) {
  if (head_dim != 128) { g_err = "ref shim is built for head_dim 128"; return 4; }
  BatchArgs a{float_ws, int_ws, pinned_ws, float_bytes, int_bytes, indptr_h, bs, hq, hkv, page_size,
              layout, cuda_graph, q, k, v, o, lse, q_stride_n, q_stride_h, {s0, s1, s2}, indices,
              indptr, last_page_len, sm_scale, static_cast<cudaStream_t>(stream), info_out, mode};
  a.rope_rcp_scale = 1.f / rope_scale;
  a.rope_rcp_theta = 1.f / rope_theta;
  try {
    if (q_dtype == 0 && kv_dtype == 0)
      return batch_impl<half, half, 128, PosEncodingMode::kRoPELlama>(a);
    if (q_dtype == 1 && kv_dtype == 1)
      return batch_impl<nv_bfloat16, nv_bfloat16, 128, PosEncodingMode::kRoPELlama>(a);
    g_err = "ROPE_LLAMA shim: 16-bit KV only";
    return 4;
  } catch (const std::exception& e) {
    g_err = e.what();
    return 2;
  }
}

// logits transforms of DefaultAttention (variants.cuh:31-93), f16 q / kv, head_dim 128:
// variant 1 = sliding window + logits soft cap, variant 2 = ALiBi
__attribute__((visibility("default"))) int ref_batch_decode_variant(
    void* float_ws, size_t float_bytes, void* int_ws, void* pinned_ws, size_t int_bytes,
    int32_t* indptr_h, int bs, int hq, int hkv, int page_size, int layout, void* q, int64_t q_stride_n,
    int64_t q_stride_h, void* k, void* v, int64_t s0, int64_t s1, int64_t s2, int32_t* indices,
    int32_t* indptr, int32_t* last_page_len, void* o, float* lse, float sm_scale, int variant,
    int window_left, float logits_soft_cap, float* alibi_slopes, void* stream) {
  BatchArgs a{float_ws, int_ws, pinned_ws, float_bytes, int_bytes, indptr_h, bs, hq, hkv, page_size,
              layout, 0, q, k, v, o, lse, q_stride_n, q_stride_h, {s0, s1, s2}, indices,
              indptr, last_page_len, sm_scale, static_cast<cudaStream_t>(stream), nullptr, 0};
  a.window_left = window_left;
  a.logits_soft_cap = logits_soft_cap;
  a.alibi_slopes = alibi_slopes;
  try {
    if (variant == 1)
      return batch_impl<half, half, 128, PosEncodingMode::kNone,
                        DefaultAttention<false, /*swa=*/true, /*softcap=*/true, false>>(a);
    if (variant == 2)
      return batch_impl<half, half, 128, PosEncodingMode::kALiBi,
                        DefaultAttention<false, false, false, /*alibi=*/true>>(a);
    g_err = "unknown variant";
    return 4;
  } catch (const std::exception& e) {
    g_err = e.what();
    return 2;
  }
}

__attribute__((visibility("default"))) int ref_single_decode(void* q, void* k, void* v, void* tmp,
                                                            void* o, float* lse, int hq, int hkv,
                                                            int head_dim, int kv_len, int q_dtype,
                                                            int kv_dtype, int layout, float sm_scale,
                                                            void* stream) {
  if (head_dim != 128) { g_err = "ref shim is built for head_dim 128"; return 4; }
  try {
    REF_DISPATCH(q_dtype, kv_dtype, return (single_impl<Q, KV, 128>(q, k, v, tmp, o, lse, hq, hkv, kv_len, layout, sm_scale, static_cast<cudaStream_t>(stream))));
  } catch (const std::exception& e) {
    g_err = e.what();
    return 2;
  }
}

#define REF_DT(dt, ...)                                       \
  if (dt == 0) { using T = half; __VA_ARGS__; }               \
  else if (dt == 1) { using T = nv_bfloat16; __VA_ARGS__; }   \
  else { g_err = "unsupported dtype"; return 4; }

__attribute__((visibility("default"))) int ref_merge_state(void* va, float* sa, void* vb, float* sb,
                                                          void* vm, float* sm, int seq, int heads,
                                                          int d, int dt, void* stream) {
  cudaError_t st;
  REF_DT(dt, st = MergeState(static_cast<T*>(va), sa, static_cast<T*>(vb), sb, static_cast<T*>(vm), sm,
                             seq, heads, d, static_cast<cudaStream_t>(stream)));
  if (st != cudaSuccess) { g_err = cudaGetErrorString(st); return 3; }
  return 0;
}

__attribute__((visibility("default"))) int ref_merge_state_in_place(void* v, float* s, void* vo,
                                                                   float* so, uint8_t* mask, int seq,
                                                                   int heads, int d, int dt,
                                                                   void* stream) {
  cudaError_t st;
  REF_DT(dt, st = MergeStateInPlace(static_cast<T*>(v), s, static_cast<T*>(vo), so, seq, heads, d, mask,
                                    static_cast<cudaStream_t>(stream)));
  if (st != cudaSuccess) { g_err = cudaGetErrorString(st); return 3; }
  return 0;
}

__attribute__((visibility("default"))) int ref_merge_states(void* v, float* s, void* vm, float* sm,
                                                           int n, int seq, int heads, int d, int dt,
                                                           void* stream) {
  cudaError_t st;
  REF_DT(dt, st = MergeStates(static_cast<T*>(v), s, static_cast<T*>(vm), sm, n, seq, heads, d,
                              static_cast<cudaStream_t>(stream)));
  if (st != cudaSuccess) { g_err = cudaGetErrorString(st); return 3; }
  return 0;
}

}  // extern "C"
