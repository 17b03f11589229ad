// decode_rope_kernel.cuh — paged decode with RoPE applied on the fly (pos_encoding_mode = "ROPE_LLAMA").
//
// The reference rotates q at position kv_len - 1 and EVERY K row at its own position inside the
// decode kernel (include/flashinfer/attention/decode.cuh:445-466 and :72-76, vec_apply_llama_rope
// in include/flashinfer/pos_enc.cuh:79-100); the cache holds un-rotated K.  A per-element rotation
// of K does not fit the TMA -> ldmatrix -> MMA path of decode_kernel.cuh (the operands never pass
// through registers as scalars there), and no serving engine uses this mode (they rotate before the
// cache write), so this variant is a plain CUDA-core kernel over the SAME planned work list:
// 4 warps per CTA, a warp takes every 4th token of a segment, a lane owns HD / 32 consecutive
// dims (its rotation partner, dim +- HD / 2, lives in lane ^ 16), fp32 math, online softmax in
// base 2.  Static plans only (one item per CTA); segment epilogue, partial slots and the fused
// split-KV merge are the ones of decode_kernel.cuh.  Not tuned: it is CUDA-core / sincos bound
// (about the reference kernel's own speed), the tuned paths are the "NONE" / "ALIBI" ones.
#pragma once
#include <cuda_fp8.h>

#include "decode_kernel.cuh"

namespace fib200 {

constexpr int kRopeThreads = 128;
constexpr int kRopeUnroll = 4;  // tokens in flight per warp

template <int NB> struct RopeRaw;
template <> struct RopeRaw<2> { using T = uint16_t; };
template <> struct RopeRaw<4> { using T = uint32_t; };
template <> struct RopeRaw<8> { using T = uint2; };
template <> struct RopeRaw<16> { using T = uint4; };

// FMT: 0 f16, 1 bf16, 2 e4m3, 3 e5m2
template <int FMT>
__device__ __forceinline__ float rope_elem(const void* raw, int i) {
  if constexpr (FMT == 0) {
    return __half2float(static_cast<const __half*>(raw)[i]);
  } else if constexpr (FMT == 1) {
    return __bfloat162float(static_cast<const __nv_bfloat16*>(raw)[i]);
  } else {
    const __nv_fp8_storage_t b = static_cast<const __nv_fp8_storage_t*>(raw)[i];
    const __half_raw h = __nv_cvt_fp8_to_halfraw(b, FMT == 2 ? __NV_E4M3 : __NV_E5M2);
    return __half2float(__half(h));
  }
}

// x[VEC] (this lane's dims) rotated by angle pos * freq[v]; partner dims come from lane ^ 16
template <int VEC>
__device__ __forceinline__ void rope_rotate(float (&x)[VEC], const float (&freq)[VEC], float pos,
                                            int lane) {
#pragma unroll
  for (int v = 0; v < VEC; ++v) {
    const float other = __shfl_xor_sync(0xffffffffu, x[v], 16);
    float sn, cs;
    sincosf(pos * freq[v], &sn, &cs);
    x[v] = x[v] * cs + (lane < 16 ? -other : other) * sn;
  }
}

// QT: 0 f16, 1 bf16 (q / o).  KVF: 0 = KV has q's dtype, 1 = e4m3, 2 = e5m2.
template <int QT, int KVF, int HD>
__global__ void __launch_bounds__(kRopeThreads)
    paged_decode_rope_kernel(const __grid_constant__ DecodeParams p) {
  constexpr int VEC = HD / 32;
  constexpr int ES = KVF == 0 ? 2 : 1;
  constexpr int KFMT = KVF == 0 ? QT : (KVF == 1 ? 2 : 3);
  using Raw = typename RopeRaw<VEC * ES>::T;
  __shared__ float sm_m[4][kMaxGroup], sm_l[4][kMaxGroup];
  __shared__ __align__(16) float sm_o[4][kMaxGroup][HD];
  __shared__ int sm_last;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int GS = p.group_size;
  const int TT = tile_tokens(KVF);
  const bool implicit = p.header == nullptr;
  const bool use_cap = p.soft_cap > 0.f;
  const float cap_pre = use_cap ? p.sm_scale / p.soft_cap : 0.f;
  const float cap_post = use_cap ? p.soft_cap * 1.4426950408889634f : 0.f;

  // ---- this CTA's segments: item blockIdx.x of the static plan, or the closed-form split
  SegSource src;
  src.p = &p; src.cache = nullptr; src.ctl = nullptr; src.implicit = implicit; src.tt = TT;
  src.first = src.head_count = src.num_items = src.count = 0;
  src.imp_begin = src.imp_end = 0;
  const ItemHead* head = nullptr;
  const Segment* list = nullptr;
  int nseg;
  if (implicit) {
    seg_source_implicit_range(src);
    nseg = src.count;
  } else {
    head = p.item_heads + blockIdx.x;
    const int4 h = __ldg(reinterpret_cast<const int4*>(head));
    src.first = h.x;
    nseg = h.y;
    list = reinterpret_cast<const Segment*>(p.item_heads + max(h.z, (int)gridDim.x));
  }
  if (nseg == 0) return;
  if (p.enable_pdl) pdl_wait();

  float freq[VEC];
#pragma unroll
  for (int v = 0; v < VEC; ++v)  // decode.cuh:455-459
    freq[v] = p.rope_rcp_scale *
              powf(p.rope_rcp_theta, (float)(2 * ((lane * VEC + v) % (HD / 2))) / (float)HD);

  const char* kb = static_cast<const char*>(p.k);
  const char* vb = static_cast<const char*>(p.v);

  for (int k = 0; k < nseg; ++k) {
    Segment sg;
    if (implicit) {
      sg = seg_make(src, k);
    } else {
      const int4* sp = reinterpret_cast<const int4*>(
          k < kItemInlineSegs ? &head->seg[k] : list + src.first + k);
      const int4 a = __ldg(sp), b = __ldg(sp + 1), c = __ldg(sp + 2);
      sg.request = a.x; sg.kv_head = a.y; sg.tile_begin = a.z; sg.tile_end = a.w;
      sg.slot = b.x; sg.tok_lo = b.y; sg.tok_hi = b.z; sg.merge_unit = b.w;
      sg.page_base = c.x; sg.merge_count = c.y; sg.merge_slot = c.z; sg.kv_len = c.w;
    }
    // ---- query heads of the group, rotated at position kv_len - 1 (decode.cuh:450,464-466)
    float qr[kMaxGroup][VEC];
#pragma unroll
    for (int j = 0; j < kMaxGroup; ++j) {
      if (j < GS) {
        const char* qrow = static_cast<const char*>(p.q) +
                           ((int64_t)sg.request * p.q_stride_n +
                            (int64_t)(sg.kv_head * GS + j) * p.q_stride_h + lane * VEC) * 2;
#pragma unroll
        for (int v = 0; v < VEC; ++v) qr[j][v] = rope_elem<QT>(qrow, v);  // q may be any strided view
        rope_rotate<VEC>(qr[j], freq, (float)(sg.kv_len - 1), lane);
      } else {
#pragma unroll
        for (int v = 0; v < VEC; ++v) qr[j][v] = 0.f;
      }
    }
    float m[kMaxGroup], l[kMaxGroup], o[kMaxGroup][VEC];
#pragma unroll
    for (int j = 0; j < kMaxGroup; ++j) {
      m[j] = kNegInf; l[j] = 0.f;
#pragma unroll
      for (int v = 0; v < VEC; ++v) o[j][v] = 0.f;
    }

    const int t_begin = max(sg.tile_begin * TT, sg.tok_lo);
    const int t_end = min(sg.tile_end * TT, sg.tok_hi);
    const int64_t head_off = (int64_t)sg.kv_head * p.kv_stride_h + lane * VEC;
    for (int base = t_begin + warp; base < t_end; base += 4 * kRopeUnroll) {
      Raw kraw[kRopeUnroll], vraw[kRopeUnroll];
#pragma unroll
      for (int u = 0; u < kRopeUnroll; ++u) {
        const int tok = base + 4 * u;
        if (tok < t_end) {  // warp-uniform
          int pit, rip;
          if (p.page_shift >= 0) { pit = tok >> p.page_shift; rip = tok & p.page_mask; }
          else { pit = tok / p.page_size; rip = tok - pit * p.page_size; }
          const int64_t page = p.kv_indices ? __ldg(p.kv_indices + sg.page_base + pit) : sg.page_base + pit;
          const int64_t off = (page * p.kv_stride_page + (int64_t)rip * p.kv_stride_n + head_off) * ES;
          kraw[u] = __ldg(reinterpret_cast<const Raw*>(kb + off));
          vraw[u] = __ldg(reinterpret_cast<const Raw*>(vb + off));
        }
      }
#pragma unroll
      for (int u = 0; u < kRopeUnroll; ++u) {
        const int tok = base + 4 * u;
        if (tok >= t_end) break;
        float kf[VEC], vf[VEC];
#pragma unroll
        for (int v = 0; v < VEC; ++v) {
          kf[v] = rope_elem<KFMT>(&kraw[u], v);
          vf[v] = rope_elem<KFMT>(&vraw[u], v);
        }
        rope_rotate<VEC>(kf, freq, (float)tok, lane);  // decode.cuh:72-76: every K row at its position
#pragma unroll
        for (int j = 0; j < kMaxGroup; ++j) {
          if (j < GS) {
            float s = 0.f;
#pragma unroll
            for (int v = 0; v < VEC; ++v) s += qr[j][v] * kf[v];
#pragma unroll
            for (int sh = 16; sh > 0; sh >>= 1) s += __shfl_xor_sync(0xffffffffu, s, sh);
            const float x = use_cap ? cap_post * tanhf_approx(s * cap_pre) : s * p.sm_scale_log2;
            const float mn = fmaxf(m[j], x);
            const float sc = ex2f(m[j] - mn);
            const float pj = ex2f(x - mn);
            l[j] = l[j] * sc + pj;
#pragma unroll
            for (int v = 0; v < VEC; ++v) o[j][v] = o[j][v] * sc + pj * vf[v];
            m[j] = mn;
          }
        }
      }
    }

    // ---- segment epilogue: 4 warp states -> one (o, lse)
    __syncthreads();  // the previous segment's combine no longer reads shared memory
#pragma unroll
    for (int j = 0; j < kMaxGroup; ++j) {
      if (j < GS) {
        if (lane == 0) { sm_m[warp][j] = m[j]; sm_l[warp][j] = l[j]; }
#pragma unroll
        for (int v = 0; v < VEC; ++v) sm_o[warp][j][lane * VEC + v] = o[j][v];
      }
    }
    __syncthreads();
    const bool direct = sg.slot < 0;
    const int64_t hq = (int64_t)sg.request * p.num_qo_heads + sg.kv_head * GS;
    for (int e = tid; e < GS * HD; e += kRopeThreads) {
      const int j = e / HD, d = e - j * HD;
      const float M = fmaxf(fmaxf(sm_m[0][j], sm_m[1][j]), fmaxf(sm_m[2][j], sm_m[3][j]));
      float L = 0.f, acc = 0.f;
#pragma unroll
      for (int w = 0; w < 4; ++w) {
        const float f = ex2f(sm_m[w][j] - M);
        L += sm_l[w][j] * f;
        acc += sm_o[w][j][d] * f;
      }
      const bool any = L > 0.f;
      const float val = any ? acc * ((direct ? p.v_scale : 1.f) / L) : 0.f;
      const float lse2 = any ? M + lg2f(L) : kNegInf;
      if (direct) {
        if (QT == 1) static_cast<__nv_bfloat16*>(p.o)[(hq + j) * HD + d] = __float2bfloat16_rn(val);
        else static_cast<__half*>(p.o)[(hq + j) * HD + d] = __float2half_rn(val);
        if (d == 0 && p.lse != nullptr) p.lse[hq + j] = lse2;
      } else {
        p.tmp_o[((int64_t)sg.slot * GS + j) * HD + d] = val;
        if (d == 0) p.tmp_lse[(int64_t)sg.slot * GS + j] = lse2;
      }
    }
    if (sg.merge_unit >= 0) {  // fused split-KV merge: the last arriver merges the unit
      __syncthreads();         // every thread's part of the partial is written
      if (tid == 0) {
        const int old = atom_add_acq_rel_gpu(p.merge_counters + sg.merge_unit, 1);
        sm_last = old == sg.merge_count - 1;
        if (sm_last) p.merge_counters[sg.merge_unit] = 0;
      }
      __syncthreads();
      if (sm_last) {
        __threadfence();
        merge_unit_cta<QT, HD>(p, sg.request, sg.kv_head, sg.merge_slot, sg.merge_count, tid);
      }
    }
  }
  if (p.enable_pdl) pdl_launch_dependents();
}

}  // namespace fib200
