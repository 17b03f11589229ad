// decode_rope_kernel.cuh — paged decode with RoPE applied on the fly (pos_encoding_mode = "ROPE_LLAMA").
//
// The reference rotates q at position kv_len - 1 and EVERY K row at its own position inside the
// decode kernel (include/flashinfer/attention/decode.cuh:445-466 and :72-76, vec_apply_llama_rope
// in include/flashinfer/pos_enc.cuh:79-100); the cache holds un-rotated K.  A per-element rotation
// of K does not fit the TMA -> ldmatrix -> MMA path of decode_kernel.cuh (the operands never pass
// through registers as scalars there), and no serving engine uses this mode (they rotate before the
// cache write), so this variant is a plain CUDA-core kernel over the SAME planned work list:
// NW = 8 warps per CTA (4 at D = 256).  A token's row is shared by 16 lanes (32 at D = 256), so a
// warp works on two tokens at once; a lane owns C = D / 32 dims in the LOW half of the head and the
// C dims D / 2 above them, i.e. both members of every rotation pair, so the rotation needs no
// shuffle and the 16 lanes read two contiguous 128-byte runs of the row.  fp32 math, online softmax
// in base 2, advanced once per U = 4 tokens (2 when the GQA group is > 4) so that the U x group dot
// products reduce through independent shuffle chains; the next U rows are loaded into registers
// before the current ones are used.  cos / sin of a lane's next token come from the previous ones
// by the constant rotation R(step f) and are re-anchored with sincosf every 32 steps.  Static plans
// only (one item per CTA); segment epilogue, partial slots and the fused split-KV merge are the
// ones of decode_kernel.cuh.  CUDA-core bound: about 1.8 TB/s at Llama-3-8B bs 64 ctx 4K (0.6 ms; the
// tuned "NONE" / "ALIBI" paths run the same shape in 0.17 ms).
#pragma once
#include <cuda_fp8.h>

#include "decode_kernel.cuh"

namespace fib200 {

constexpr int kRopeAnchor = 32;  // steps between exact sincosf evaluations
__host__ __device__ constexpr int rope_warps(int hd) { return hd <= 128 ? 8 : 4; }
// lanes that share one token: 16 up to D = 128 (a warp works on two tokens), the whole warp at 256
__host__ __device__ constexpr int rope_lanes_per_token(int hd) { return hd <= 128 ? 16 : 32; }
// tokens a lane group keeps in flight (registers hold this many K / V rows twice: current + next)
__host__ __device__ constexpr int rope_unroll(int gb, int hd = 128) { return hd >= 512 ? 1 : (gb <= 4 ? 4 : 2); }

template <int NB> struct RopeRaw;
template <> struct RopeRaw<2> { using T = uint16_t; static __device__ __forceinline__ uint32_t word(T r, int) { return r; } };
template <> struct RopeRaw<4> { using T = uint32_t; static __device__ __forceinline__ uint32_t word(T r, int) { return r; } };
template <> struct RopeRaw<8> { using T = uint2; static __device__ __forceinline__ uint32_t word(T r, int i) { return i == 0 ? r.x : r.y; } };
template <> struct RopeRaw<16> { using T = uint4; static __device__ __forceinline__ uint32_t word(T r, int i) { return i == 0 ? r.x : (i == 1 ? r.y : (i == 2 ? r.z : r.w)); } };

// elements 2i, 2i + 1 of a packed row fragment.  FMT: 0 f16, 1 bf16, 2 e4m3, 3 e5m2
template <int FMT, int NB>
__device__ __forceinline__ float2 rope_pair(typename RopeRaw<NB>::T r, int i) {
  if constexpr (FMT == 0) {
    const uint32_t w = RopeRaw<NB>::word(r, i);
    return __half22float2(*reinterpret_cast<const __half2*>(&w));
  } else if constexpr (FMT == 1) {
    const uint32_t w = RopeRaw<NB>::word(r, i);
    return make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xffff0000u));
  } else {
    const uint32_t w = RopeRaw<NB>::word(r, i >> 1) >> (16 * (i & 1));
    const __half2_raw h = __nv_cvt_fp8x2_to_halfraw2((__nv_fp8x2_storage_t)(w & 0xffffu),
                                                     FMT == 2 ? __NV_E4M3 : __NV_E5M2);
    return __half22float2(__half2(h));
  }
}

__device__ __forceinline__ float rope_q_elem(const void* row, int i, int qt) {
  return qt == 0 ? __half2float(static_cast<const __half*>(row)[i])
                 : __bfloat162float(static_cast<const __nv_bfloat16*>(row)[i]);
}

// QT: 0 f16, 1 bf16 (q / o).  KVF: 0 = KV has q's dtype, 1 = e4m3, 2 = e5m2.  GB: compile-time
// bound on the GQA group (4 or 8); heads past group_size carry q = 0 and are never stored.
template <int QT, int KVF, int HD, int GB>
__global__ void __launch_bounds__(rope_warps(HD) * 32)
    paged_decode_rope_kernel(const __grid_constant__ DecodeParams p) {
  constexpr int NW = rope_warps(HD);
  constexpr int LPT = rope_lanes_per_token(HD);
  constexpr int TPW = 32 / LPT;        // tokens a warp works on at once
  constexpr int C = HD / (2 * LPT);    // dims a lane owns in EACH half of the head
  constexpr int U = rope_unroll(GB, HD);
  constexpr int STEP = NW * TPW;       // tokens the CTA covers per unrolled step
  constexpr int ES = KVF == 0 ? 2 : 1;
  constexpr int KFMT = KVF == 0 ? QT : (KVF == 1 ? 2 : 3);
  using RR = RopeRaw<C * ES>;
  using Raw = typename RR::T;
  __shared__ float sm_m[NW][kMaxGroup], sm_l[NW][kMaxGroup];
  __shared__ __align__(16) float sm_o[NW][GB][HD];  // 32 KiB at most (HD 256 x 8 heads, HD 512 x 4 heads)
  __shared__ int sm_last;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int sub = lane & (LPT - 1), half = lane / LPT;
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
  if (p.enable_pdl == 1) pdl_launch_dependents();  // early trigger, see paged_decode_kernel

  // dims (C sub + c) and (HD / 2 + C sub + c) are a rotation pair and share one frequency
  float freq[C], cstep[C], sstep[C];
#pragma unroll
  for (int c = 0; c < C; ++c) {  // decode.cuh:455-459
    freq[c] = p.rope_rcp_scale * powf(p.rope_rcp_theta, (float)(2 * (sub * C + c)) / (float)HD);
    sincosf((float)STEP * freq[c], &sstep[c], &cstep[c]);
  }

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
    float qlo[GB][C], qhi[GB][C];
    {
      float qc[C], qs[C];
#pragma unroll
      for (int c = 0; c < C; ++c) sincosf((float)(sg.kv_len - 1) * freq[c], &qs[c], &qc[c]);
#pragma unroll
      for (int j = 0; j < GB; ++j) {
        if (j < GS) {
          const char* qrow = static_cast<const char*>(p.q) +
                             ((int64_t)sg.request * p.q_stride_n +
                              (int64_t)(sg.kv_head * GS + j) * p.q_stride_h) * 2;
#pragma unroll
          for (int c = 0; c < C; ++c) {  // q may be any strided view: scalar loads
            const float a = rope_q_elem(qrow, sub * C + c, QT);
            const float b = rope_q_elem(qrow, HD / 2 + sub * C + c, QT);
            qlo[j][c] = a * qc[c] - b * qs[c];
            qhi[j][c] = b * qc[c] + a * qs[c];
          }
        } else {
#pragma unroll
          for (int c = 0; c < C; ++c) qlo[j][c] = qhi[j][c] = 0.f;
        }
      }
    }
    float m[GB], l[GB], olo[GB][C], ohi[GB][C];
#pragma unroll
    for (int j = 0; j < GB; ++j) {
      m[j] = kNegInf; l[j] = 0.f;
#pragma unroll
      for (int c = 0; c < C; ++c) olo[j][c] = ohi[j][c] = 0.f;
    }

    const int t_begin = max(sg.tile_begin * TT, sg.tok_lo);
    const int t_end = min(sg.tile_end * TT, sg.tok_hi);
    const int64_t head_off = (int64_t)(sg.kv_head / p.kv_head_div) * p.kv_stride_h + sub * C;

    // rows of U tokens: [u][0] = the low-half fragment, [u][1] = its rotation partners
    auto load_rows = [&](int first, Raw (&kr)[U][2], Raw (&vr)[U][2]) {
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int tok = first + u * STEP;
        if (tok < t_end) {
          int pit, rip;
          if (p.page_shift >= 0) { pit = tok >> p.page_shift; rip = tok & p.page_mask; }
          else { pit = tok / p.page_size; rip = tok - pit * p.page_size; }
          const int64_t page = p.kv_indices ? __ldg(p.kv_indices + sg.page_base + pit) : sg.page_base + pit;
          const int64_t off = (page * p.kv_stride_page + (int64_t)rip * p.kv_stride_n + head_off) * ES;
          kr[u][0] = __ldg(reinterpret_cast<const Raw*>(kb + off));
          kr[u][1] = __ldg(reinterpret_cast<const Raw*>(kb + off + (HD / 2) * ES));
          vr[u][0] = __ldg(reinterpret_cast<const Raw*>(vb + off));
          vr[u][1] = __ldg(reinterpret_cast<const Raw*>(vb + off + (HD / 2) * ES));
        } else {  // zero rows: their weight is forced to 0 below, 0 x garbage must not be NaN
          kr[u][0] = kr[u][1] = vr[u][0] = vr[u][1] = Raw{};
        }
      }
    };

    float cs[C], sn[C];
    int since_anchor = 0;
    Raw kcur[U][2], vcur[U][2];
    const int my_first = t_begin + warp * TPW + half;
    load_rows(my_first, kcur, vcur);
    // the loop bound is the warp's first token (warp-uniform: full-mask shuffles inside)
    for (int wbase = t_begin + warp * TPW; wbase < t_end; wbase += U * STEP) {
      const int first = wbase + half;
      Raw knext[U][2], vnext[U][2];
      if (wbase + U * STEP < t_end) load_rows(first + U * STEP, knext, vnext);
      float x[U][GB], vlo[U][C], vhi[U][C];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int tok = first + u * STEP;
        float klo[C], khi[C];
#pragma unroll
        for (int c2 = 0; c2 < C / 2; ++c2) {
          const float2 a = rope_pair<KFMT, C * ES>(kcur[u][0], c2), b = rope_pair<KFMT, C * ES>(kcur[u][1], c2);
          const float2 d = rope_pair<KFMT, C * ES>(vcur[u][0], c2), e = rope_pair<KFMT, C * ES>(vcur[u][1], c2);
          klo[2 * c2] = a.x; klo[2 * c2 + 1] = a.y; khi[2 * c2] = b.x; khi[2 * c2 + 1] = b.y;
          vlo[u][2 * c2] = d.x; vlo[u][2 * c2 + 1] = d.y; vhi[u][2 * c2] = e.x; vhi[u][2 * c2 + 1] = e.y;
        }
        // decode.cuh:72-76: every K row rotated at its own position
        if (since_anchor == 0) {
#pragma unroll
          for (int c = 0; c < C; ++c) sincosf((float)tok * freq[c], &sn[c], &cs[c]);
        }
        if (++since_anchor == kRopeAnchor) since_anchor = 0;
#pragma unroll
        for (int c = 0; c < C; ++c) {
          const float a = klo[c], b = khi[c];
          klo[c] = a * cs[c] - b * sn[c];
          khi[c] = b * cs[c] + a * sn[c];
          const float cn = cs[c] * cstep[c] - sn[c] * sstep[c];  // this lane's next token: + STEP
          sn[c] = sn[c] * cstep[c] + cs[c] * sstep[c];
          cs[c] = cn;
        }
#pragma unroll
        for (int j = 0; j < GB; ++j) {
          float s = 0.f;
#pragma unroll
          for (int c = 0; c < C; ++c) s += qlo[j][c] * klo[c] + qhi[j][c] * khi[c];
          x[u][j] = s;
        }
      }
      // the U x GB partial dot products: butterfly over the token's LPT lanes (independent chains)
#pragma unroll
      for (int sh = LPT / 2; sh > 0; sh >>= 1) {
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
          for (int j = 0; j < GB; ++j) x[u][j] += __shfl_xor_sync(0xffffffffu, x[u][j], sh);
      }
      // one online-softmax step for the U tokens of every head
#pragma unroll
      for (int j = 0; j < GB; ++j) {
        float mn = m[j];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          x[u][j] = use_cap ? cap_post * tanhf_approx(x[u][j] * cap_pre) : x[u][j] * p.sm_scale_log2;
          if (first + u * STEP < t_end) mn = fmaxf(mn, x[u][j]);
        }
        const float sc = ex2f(m[j] - mn);
        m[j] = mn;
        l[j] *= sc;
#pragma unroll
        for (int c = 0; c < C; ++c) { olo[j][c] *= sc; ohi[j][c] *= sc; }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const float pj = first + u * STEP < t_end ? ex2f(x[u][j] - mn) : 0.f;
          l[j] += pj;
#pragma unroll
          for (int c = 0; c < C; ++c) { olo[j][c] += pj * vlo[u][c]; ohi[j][c] += pj * vhi[u][c]; }
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        kcur[u][0] = knext[u][0]; kcur[u][1] = knext[u][1];
        vcur[u][0] = vnext[u][0]; vcur[u][1] = vnext[u][1];
      }
    }

    // ---- segment epilogue: (two lane groups per warp ->) NW warp states -> one (o, lse)
    if constexpr (TPW == 2) {
#pragma unroll
      for (int j = 0; j < GB; ++j) {
        const float mo = __shfl_xor_sync(0xffffffffu, m[j], 16);
        const float lo = __shfl_xor_sync(0xffffffffu, l[j], 16);
        const float M = fmaxf(m[j], mo);
        const float fa = ex2f(m[j] - M), fb = ex2f(mo - M);
        m[j] = M;
        l[j] = l[j] * fa + lo * fb;
#pragma unroll
        for (int c = 0; c < C; ++c) {
          olo[j][c] = olo[j][c] * fa + __shfl_xor_sync(0xffffffffu, olo[j][c], 16) * fb;
          ohi[j][c] = ohi[j][c] * fa + __shfl_xor_sync(0xffffffffu, ohi[j][c], 16) * fb;
        }
      }
    }
    __syncthreads();  // the previous segment's combine no longer reads shared memory
    if (half == 0) {
#pragma unroll
      for (int j = 0; j < GB; ++j) {
        if (j < GS) {
          if (sub == 0) { sm_m[warp][j] = m[j]; sm_l[warp][j] = l[j]; }
#pragma unroll
          for (int c = 0; c < C; ++c) {
            sm_o[warp][j][sub * C + c] = olo[j][c];
            sm_o[warp][j][HD / 2 + sub * C + c] = ohi[j][c];
          }
        }
      }
    }
    __syncthreads();
    const bool direct = sg.slot < 0;
    const int64_t hq = (int64_t)sg.request * p.num_qo_heads + sg.kv_head * GS;
    for (int e = tid; e < GS * HD; e += NW * 32) {
      const int j = e / HD, d = e - j * HD;
      float M = sm_m[0][j];
#pragma unroll
      for (int w = 1; w < NW; ++w) M = fmaxf(M, sm_m[w][j]);
      float L = 0.f, acc = 0.f;
#pragma unroll
      for (int w = 0; w < NW; ++w) {
        const float f = ex2f(sm_m[w][j] - M);
        L += sm_l[w][j] * f;
        acc += sm_o[w][j][d] * f;
      }
      const bool any = L > 0.f;
      if (direct && any && p.sinks != nullptr)
        L += ex2f(__ldg(p.sinks + sg.kv_head * GS + j) * 1.4426950408889634f - M);
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
      if (sm_last && tid < kConsumerWarps * 32) {  // merge_unit_cta strides over 128 threads
        __threadfence();
        merge_unit_cta<QT, HD>(p, sg.request, sg.kv_head, sg.merge_slot, sg.merge_count, tid);
      }
    }
  }
  if (p.enable_pdl == 2) pdl_launch_dependents();
}

}  // namespace fib200
