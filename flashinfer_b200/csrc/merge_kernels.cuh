// merge_kernels.cuh — attention-state merge kernels (base-2 LSE).
//
// Replaces the reference's cascade kernels (include/flashinfer/attention/cascade.cuh):
//   MergeStateKernel :44-71, MergeStateInPlaceKernel :86-116, MergeStatesKernel :213-256,
//   MergeStatesLargeNumIndexSetsKernel :275-340,
//   PersistentVariableLengthMergeStatesKernel :366-467.
// The merge operator: given states (v_i, s_i) with s_i = log2-sum-exp,
//   s = log2(sum_i 2^{s_i}),  v = sum_i 2^{s_i - s} v_i      (reference state.cuh:53-62)
// All of them are pure streaming kernels (HBM-bound): every thread owns 8 contiguous
// head-dim elements (16-byte loads for 16-bit inputs), fp32 math, ex2/lg2.approx.
#pragma once
#include "decode_params.h"
#include "ptx.cuh"

namespace fib200 {

// DT: 0 f16, 1 bf16, 4 f32
template <int DT>
struct Vec8 {
  float x[8];
  __device__ __forceinline__ void load(const void* base, int64_t elem_off) {
    if constexpr (DT == 4) {
      const float4* p = reinterpret_cast<const float4*>(static_cast<const float*>(base) + elem_off);
      float4 a = p[0], b = p[1];
      x[0] = a.x; x[1] = a.y; x[2] = a.z; x[3] = a.w;
      x[4] = b.x; x[5] = b.y; x[6] = b.z; x[7] = b.w;
    } else {
      const uint4 u = *reinterpret_cast<const uint4*>(static_cast<const uint16_t*>(base) + elem_off);
      const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        if constexpr (DT == 1) {
          x[2 * i] = __uint_as_float(w[i] << 16);
          x[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
        } else {
          const __half2 h = *reinterpret_cast<const __half2*>(&w[i]);
          const float2 f = __half22float2(h);
          x[2 * i] = f.x;
          x[2 * i + 1] = f.y;
        }
      }
    }
  }
  __device__ __forceinline__ void store(void* base, int64_t elem_off) const {
    if constexpr (DT == 4) {
      float4* p = reinterpret_cast<float4*>(static_cast<float*>(base) + elem_off);
      p[0] = make_float4(x[0], x[1], x[2], x[3]);
      p[1] = make_float4(x[4], x[5], x[6], x[7]);
    } else {
      uint4 u;
      u.x = pack2<DT == 1>(x[0], x[1]);
      u.y = pack2<DT == 1>(x[2], x[3]);
      u.z = pack2<DT == 1>(x[4], x[5]);
      u.w = pack2<DT == 1>(x[6], x[7]);
      *reinterpret_cast<uint4*>(static_cast<uint16_t*>(base) + elem_off) = u;
    }
  }
};

// ---- 2-way merge: one thread per (row = seq*H + head, 8 dims) -----------------
// IN_PLACE: (v_a, s_a) is updated; rows with mask[seq] == 0 are skipped.
template <int DT, bool IN_PLACE>
__global__ void merge_state_kernel(const void* v_a, const float* s_a, const void* v_b,
                                   const float* s_b, void* v_out, float* s_out,
                                   const uint8_t* mask, int64_t rows, int num_heads, int head_dim) {
  const int vpr = head_dim / 8;  // vectors per row
  const int64_t total = rows * vpr;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t row = i / vpr;
    const int c = (int)(i - row * vpr);
    if (IN_PLACE && mask != nullptr && mask[row / num_heads] == 0) continue;
    const float sa = s_a[row], sb = s_b[row];
    const float mx = fmaxf(sa, sb);
    const float a = ex2f(sa - mx), b = ex2f(sb - mx);
    const float inv = 1.f / (a + b);
    Vec8<DT> va, vb;
    va.load(v_a, row * head_dim + c * 8);
    vb.load(v_b, row * head_dim + c * 8);
#pragma unroll
    for (int j = 0; j < 8; ++j) va.x[j] = (a * va.x[j] + b * vb.x[j]) * inv;
    va.store(v_out, row * head_dim + c * 8);
    if (c == 0) s_out[row] = lg2f(a + b) + mx;
  }
}

// ---- n-way merge over v[seq, n, H, D], s[seq, n, H] ---------------------------
// One thread per (seq, head, 8 dims); the n states are visited in index order.
template <int DT>
__global__ void merge_states_kernel(const void* v, const float* s, void* v_out, float* s_out,
                                    int64_t seq_len, int n, int num_heads, int head_dim) {
  const int vpr = head_dim / 8;
  const int64_t total = seq_len * num_heads * vpr;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % vpr);
    const int64_t row = i / vpr;  // seq * H + head
    const int64_t pos = row / num_heads;
    const int head = (int)(row - pos * num_heads);
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float m = kNegInf, d = 0.f;
    if (n == 0) { m = kNegInf; d = 1.f; }
    for (int k = 0; k < n; ++k) {
      const int64_t r = (pos * n + k) * num_heads + head;
      const float sk = s[r];
      Vec8<DT> x;
      x.load(v, r * head_dim + c * 8);
      const float mn = fmaxf(m, sk);
      const float so = ex2f(m - mn), sn = ex2f(sk - mn);
      d = d * so + sn;
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = acc[j] * so + x.x[j] * sn;
      m = mn;
    }
    const float inv = d > 0.f ? 1.f / d : 0.f;
    Vec8<DT> out;
#pragma unroll
    for (int j = 0; j < 8; ++j) out.x[j] = acc[j] * inv;
    out.store(v_out, row * head_dim + c * 8);
    if (c == 0) s_out[row] = m + lg2f(d);
  }
}

// ---- variable-length merge: v[total, H, D], indptr[seq+1] -> v_out[seq, H, D] ---
template <int DT_IN, int DT_OUT>
__global__ void varlen_merge_states_kernel(const void* v, const float* s, const int32_t* indptr,
                                           void* v_out, float* s_out, int64_t seq_len,
                                           int num_heads, int head_dim, int skip_empty,
                                           float out_scale) {
  const int vpr = head_dim / 8;
  const int64_t total = seq_len * num_heads * vpr;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % vpr);
    const int64_t row = i / vpr;
    const int64_t pos = row / num_heads;
    const int head = (int)(row - pos * num_heads);
    const int b = indptr[pos], e = indptr[pos + 1];
    if (e == b && skip_empty) continue;
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float m = kNegInf, d = (e == b) ? 1.f : 0.f;
    for (int k = b; k < e; ++k) {
      const int64_t r = (int64_t)k * num_heads + head;
      const float sk = s[r];
      Vec8<DT_IN> x;
      x.load(v, r * head_dim + c * 8);
      const float mn = fmaxf(m, sk);
      const float so = ex2f(m - mn), sn = ex2f(sk - mn);
      d = d * so + sn;
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = acc[j] * so + x.x[j] * sn;
      m = mn;
    }
    const float inv = d > 0.f ? out_scale / d : 0.f;
    Vec8<DT_OUT> out;
#pragma unroll
    for (int j = 0; j < 8; ++j) out.x[j] = acc[j] * inv;
    out.store(v_out, row * head_dim + c * 8);
    if (c == 0 && s_out != nullptr) s_out[row] = m + lg2f(d);
  }
}

// ---- split-KV merge after the decode kernel ------------------------------------
// One CTA per split (request, kv_head) unit; thread = (head in group, 4 dims).
// Partials: tmp_o[slot, G, D] fp32 (normalised), tmp_lse[slot, G].
template <int DT_OUT>
__global__ void decode_merge_kernel(const __grid_constant__ MergeParams p) {
  if (p.enable_pdl) pdl_wait();
  int request, kv_head, slot_begin, count;
  if (p.header != nullptr) {
    if ((int)blockIdx.x >= p.header->num_merge_units) return;
    const int4 u = *reinterpret_cast<const int4*>(p.units + blockIdx.x);
    request = u.x; kv_head = u.y; slot_begin = u.z; count = u.w;
  } else {
    // implicit uniform split (see seg_make): unit u spans stream [u nt, (u+1) nt); the CTA
    // holding stream position x is ((x+1) C - 1) / T; slot of CTA i's piece = i + u.
    const long long nt = p.imp_tiles, C = p.imp_ctas, T = nt * p.imp_units;
    const long long u = blockIdx.x;
    if (u >= p.imp_units || T == 0) return;
    const long long first = ((u * nt + 1) * C - 1) / T;
    const long long last = (((u + 1) * nt) * C - 1) / T;
    if (first == last) return;  // unit handled whole by one CTA
    request = 0; kv_head = (int)u; slot_begin = (int)(first + u); count = (int)(last - first + 1);
  }
  const int G = p.group_size, D = p.head_dim;
  const int vpr = D / 4;
  for (int i = threadIdx.x; i < G * vpr; i += blockDim.x) {
    const int head = i / vpr, c = i - head * vpr;
    float m = kNegInf;
    for (int k = 0; k < count; ++k) m = fmaxf(m, p.tmp_lse[(int64_t)(slot_begin + k) * G + head]);
    float d = 0.f;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int k = 0; k < count; ++k) {
      const int64_t r = (int64_t)(slot_begin + k) * G + head;
      const float w = ex2f(p.tmp_lse[r] - m);
      const float4 x = *reinterpret_cast<const float4*>(p.tmp_o + r * D + c * 4);
      d += w;
      acc.x += w * x.x; acc.y += w * x.y; acc.z += w * x.z; acc.w += w * x.w;
    }
    // partials of empty pieces carry lse = -5e4 and o = 0; if every piece is empty the
    // weights are all 1 and o stays 0, lse = -5e4 + log2(count): clamp to the sentinel
    const bool any = m > kNegInf;
    if (any && p.sinks != nullptr) d += ex2f(__ldg(p.sinks + kv_head * G + head) * 1.4426950408889634f - m);
    const float inv = p.out_scale / d;
    const int64_t orow = ((int64_t)request * p.num_qo_heads + kv_head * G + head) * D + c * 4;
    if constexpr (DT_OUT == 1) {
      uint2 u;
      u.x = pack2<true>(acc.x * inv, acc.y * inv);
      u.y = pack2<true>(acc.z * inv, acc.w * inv);
      *reinterpret_cast<uint2*>(static_cast<uint16_t*>(p.o) + orow) = u;
    } else {
      uint2 u;
      u.x = pack2<false>(acc.x * inv, acc.y * inv);
      u.y = pack2<false>(acc.z * inv, acc.w * inv);
      *reinterpret_cast<uint2*>(static_cast<uint16_t*>(p.o) + orow) = u;
    }
    if (c == 0 && p.lse != nullptr)
      p.lse[(int64_t)request * p.num_qo_heads + kv_head * G + head] = any ? m + lg2f(d) : kNegInf;
  }
}

}  // namespace fib200
