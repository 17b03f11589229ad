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

#include "decode_params.h"
#include "ptx.cuh"

namespace fib200 {

template <int KVF, int HD>
struct TileGeom {
  static constexpr int ES = (KVF == 0) ? 2 : 1;          // bytes per KV element
  static constexpr int ROW_BYTES = HD * ES;              // one token row of K (or V)
  static constexpr int NP = ROW_BYTES / 128;             // 128-byte swizzle panels per row
  static constexpr int CHUNKS = ROW_BYTES / 16;          // 16-byte chunks per row
  static constexpr int PANEL_BYTES = kTileTokens * 128;  // 8 KiB
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
  static constexpr int SCRATCH_ML = kConsumerWarps * kMaxGroup * 2 * 4;
  static constexpr int SEG_BYTES = kSegCache * (int)sizeof(Segment);
  static constexpr int BAR_BYTES = 2 * kMaxStages * 8;
  __host__ __device__ static constexpr int ring(int stages) { return stages * G::STAGE_BYTES; }
  __host__ __device__ static constexpr int total(int stages) {
    return 1024 + ring(stages) + SCRATCH_O + SCRATCH_ML + SEG_BYTES + BAR_BYTES;
  }
};

// byte offset of (row, 16B-chunk) inside the K (or V) part of a stage, 128B swizzle
__device__ __forceinline__ uint32_t swz(int row, int chunk) {
  return (uint32_t)((chunk >> 3) * (kTileTokens * 128) + row * 128 +
                    (((chunk & 7) ^ (row & 7)) << 4));
}

// ---- segment source: explicit work list or closed-form uniform split --------
struct SegSource {
  const DecodeParams* p;
  int first;  // explicit: index of this CTA's first segment
  int count;
  long long imp_begin, imp_end;  // implicit: stream range of this CTA
};

__device__ __forceinline__ SegSource seg_source_init(const DecodeParams& p) {
  SegSource s;
  s.p = &p;
  s.first = 0;
  s.count = 0;
  s.imp_begin = s.imp_end = 0;
  if (p.header != nullptr) {
    if ((int)blockIdx.x < p.header->num_ctas_used) {
      s.first = p.cta_seg_indptr[blockIdx.x];
      s.count = p.cta_seg_indptr[blockIdx.x + 1] - s.first;
    }
  } else {
    const long long T = (long long)p.imp_units * p.imp_tiles;
    const long long C = gridDim.x;
    if (T == 0) {
      s.count = (blockIdx.x == 0) ? p.imp_units : 0;  // empty KV: zero-tile segments
    } else {
      s.imp_begin = ((long long)blockIdx.x * T) / C;
      s.imp_end = ((long long)(blockIdx.x + 1) * T) / C;
      if (s.imp_end > s.imp_begin)
        s.count = (int)((s.imp_end - 1) / p.imp_tiles - s.imp_begin / p.imp_tiles) + 1;
    }
  }
  return s;
}

__device__ __forceinline__ Segment seg_make(const SegSource& s, int k) {
  const DecodeParams& p = *s.p;
  Segment g;
  if (p.header != nullptr) {
    const int4* src = reinterpret_cast<const int4*>(p.segments + s.first + k);
    int4 a = __ldg(src), b = __ldg(src + 1);
    g.request = a.x; g.kv_head = a.y; g.tile_begin = a.z; g.tile_end = a.w;
    g.slot = b.x; g.tok_lo = b.y; g.tok_hi = b.z; g.pad = 0;
  } else {
    g.request = 0; g.pad = 0;
    g.tok_lo = p.imp_tok_lo; g.tok_hi = p.imp_tok_hi;
    const int nt = p.imp_tiles;
    if (nt == 0) {
      g.kv_head = k; g.tile_begin = g.tile_end = 0; g.slot = -1;
    } else {
      const int tb0 = p.imp_tok_lo / kTileTokens;  // first tile of the window
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
  }
  return g;
}

__device__ __forceinline__ Segment seg_get(const SegSource& s, const Segment* cache, int k) {
  if (k < kSegCache) return cache[k];
  return seg_make(s, k);
}

// ---- producer: per-page TMA boxes -------------------------------------------
// Runs in one warp (all 32 lanes converged).  Box i of a tile is issued by lane i%32:
//   i = (kv * NP + panel) * BPT + j,  j = token-block of rows_per_box rows.
template <int KVF, int HD>
__device__ __forceinline__ void produce_tma(const DecodeParams& p, const SegSource& src,
                                            const Segment* seg_cache, const CUtensorMap* kmap,
                                            const CUtensorMap* vmap, uint32_t ring,
                                            uint32_t full_bar, uint32_t empty_bar) {
  using G = TileGeom<KVF, HD>;
  const int lane = threadIdx.x & 31;
  const int rpb = p.rows_per_box;
  const int bpt = kTileTokens / rpb;  // boxes per tile along tokens: 1, 2, 4 or 8
  const int nbox = 2 * G::NP * bpt;
  const int j = lane % bpt;  // same j for every i = lane + 32 k since bpt | 32
  const uint32_t box_bytes = (uint32_t)rpb * 128u;
  int stage = 0;
  uint32_t phase = 0;
  for (int k = 0; k < src.count; ++k) {
    const Segment sg = seg_get(src, seg_cache, k);
    const int page_base = p.kv_indptr ? __ldg(p.kv_indptr + sg.request) : 0;
    for (int t = sg.tile_begin; t < sg.tile_end; ++t) {
      const int tok0 = t * kTileTokens;
      const int rem = sg.tok_hi - tok0;  // > 0 by construction
      const int nvb = rem >= kTileTokens ? bpt : (rem + rpb - 1) / rpb;
      const int tok = tok0 + j * rpb;
      int page = 0;
      if (j < nvb) {
        const int pit = tok >> p.page_shift;
        page = p.kv_indices ? __ldg(p.kv_indices + page_base + pit) : page_base + pit;
      }
      const int row0 = tok & p.page_mask;
      mbar_wait(empty_bar + 8 * stage, phase ^ 1);
      if (lane == 0) mbar_arrive_expect_tx(full_bar + 8 * stage, 2u * G::NP * nvb * box_bytes);
      __syncwarp();
      if (j < nvb) {
        const uint32_t sbase = ring + stage * G::STAGE_BYTES + j * box_bytes;
        for (int i = lane; i < nbox; i += 32) {
          const int pn = (i / bpt) % G::NP;
          const int kv = i / (bpt * G::NP);
          tma_load_4d(sbase + kv * G::KV_BYTES + pn * G::PANEL_BYTES, kv ? vmap : kmap,
                      full_bar + 8 * stage, pn * (128 / G::ES), p.tma_h_first ? sg.kv_head : row0,
                      p.tma_h_first ? row0 : sg.kv_head, page);
        }
      }
      __syncwarp();
      if (++stage == p.num_stages) { stage = 0; phase ^= 1; }
    }
  }
}

// ---- producer: 16-byte cp.async gathers (any page size / stride) -------------
// Runs in 4 warps (128 threads, ptid = 0..127).  Rows past tok_hi are zero-filled.
template <int KVF, int HD>
__device__ __forceinline__ void produce_cpasync(const DecodeParams& p, const SegSource& src,
                                                const Segment* seg_cache, int ptid,
                                                uint32_t ring, uint32_t full_bar,
                                                uint32_t empty_bar) {
  using G = TileGeom<KVF, HD>;
  constexpr int CPR = G::CHUNKS;
  constexpr int RPP = 128 / CPR;             // rows per pass
  constexpr int PASSES = kTileTokens / RPP;
  static_assert(128 % CPR == 0 && kTileTokens % RPP == 0, "bad cp.async geometry");
  const int c = ptid % CPR;
  const int r0 = ptid / CPR;
  const char* kb = static_cast<const char*>(p.k);
  const char* vb = static_cast<const char*>(p.v);
  int stage = 0;
  uint32_t phase = 0;
  for (int k = 0; k < src.count; ++k) {
    const Segment sg = seg_get(src, seg_cache, k);
    const int page_base = p.kv_indptr ? __ldg(p.kv_indptr + sg.request) : 0;
    const int64_t head_off = (int64_t)sg.kv_head * p.kv_stride_h + c * (16 / G::ES);
    for (int t = sg.tile_begin; t < sg.tile_end; ++t) {
      mbar_wait(empty_bar + 8 * stage, phase ^ 1);
      const uint32_t sk = ring + stage * G::STAGE_BYTES;
#pragma unroll 4
      for (int ps = 0; ps < PASSES; ++ps) {
        const int row = ps * RPP + r0;
        const int tok = t * kTileTokens + row;
        const bool valid = tok < sg.tok_hi;
        int64_t off = 0;
        if (valid) {
          int pit, rip;
          if (p.page_shift >= 0) { pit = tok >> p.page_shift; rip = tok & p.page_mask; }
          else { pit = tok / p.page_size; rip = tok - pit * p.page_size; }
          const int page = p.kv_indices ? __ldg(p.kv_indices + page_base + pit) : page_base + pit;
          off = ((int64_t)page * p.kv_stride_page + (int64_t)rip * p.kv_stride_n + head_off) *
                G::ES;
        }
        const uint32_t d = sk + swz(row, c);
        cp_async_16(d, kb + off, valid ? 16u : 0u);
        cp_async_16(d + G::KV_BYTES, vb + off, valid ? 16u : 0u);
      }
      cp_async_mbar_arrive_noinc(full_bar + 8 * stage);
      if (++stage == p.num_stages) { stage = 0; phase ^= 1; }
    }
  }
}

// ---- consumer helpers ---------------------------------------------------------
// Query fragments (B operand of S^T = K Q^T): lane (g = lane/4, t = lane%4) holds head g.
//   16-bit KV: k-step ks covers d in [16ks, 16ks+16): b0 = q[16ks+2t..+1], b1 = q[16ks+8+2t..+1]
//   fp8 KV   : k-step c  covers d in [16c, 16c+16) with the k-slots permuted to match the
//              4-byte fp8 groups ldmatrix hands out: b0 = q[16c+4t..+1], b1 = q[16c+4t+2..+3]
template <int QT, int KVF, int HD>
__device__ __forceinline__ void load_q_frags(uint32_t (&qf)[HD / 16][2], const void* qrow,
                                             bool head_valid, int t) {
  constexpr int KS = HD / 16;
  if (!head_valid) {
#pragma unroll
    for (int i = 0; i < KS; ++i) qf[i][0] = qf[i][1] = 0u;
    return;
  }
  if constexpr (KVF == 0) {
    const uint32_t* q32 = static_cast<const uint32_t*>(qrow);
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
      qf[ks][0] = __ldg(q32 + 8 * ks + t);
      qf[ks][1] = __ldg(q32 + 8 * ks + 4 + t);
    }
  } else {
    const uint2* q64 = static_cast<const uint2*>(qrow);
#pragma unroll
    for (int c = 0; c < KS; ++c) {
      uint2 v = __ldg(q64 + 4 * c + t);  // q[16c + 4t .. 16c + 4t + 3]
      if constexpr (QT == 1) {           // bf16 -> f16 (MMA runs in f16 for fp8 KV)
        float a = __uint_as_float(v.x << 16), b = __uint_as_float(v.x & 0xffff0000u);
        float cc = __uint_as_float(v.y << 16), d = __uint_as_float(v.y & 0xffff0000u);
        v.x = pack2<false>(a, b);
        v.y = pack2<false>(cc, d);
      }
      qf[c][0] = v.x;
      qf[c][1] = v.y;
    }
  }
}

// per-lane row/chunk offsets of the ldmatrix.x4 address patterns
//   K (non-trans) and fp8 V : row = (lane&7) + 8*((lane>>3)&1), chunk += lane>>4
//   16-bit V (trans)         : row = (lane&7) + 8*(lane>>4),     chunk += (lane>>3)&1
struct LaneAddr {
  int krow, kch;  // K pattern
  int vrow, vch;  // V pattern
};
template <int KVF>
__device__ __forceinline__ LaneAddr lane_addr(int warp, int lane) {
  LaneAddr a;
  a.krow = warp * 16 + (lane & 7) + 8 * ((lane >> 3) & 1);
  a.kch = lane >> 4;
  if constexpr (KVF == 0) {
    a.vrow = warp * 16 + (lane & 7) + 8 * (lane >> 4);
    a.vch = (lane >> 3) & 1;
  } else {
    a.vrow = a.krow;
    a.vch = a.kch;
  }
  return a;
}

// S^T(16 tokens x 8 heads) for this warp's token slice
template <int QT, int KVF, int HD>
__device__ __forceinline__ void qk_tile(float (&s)[4], uint32_t kbase, const LaneAddr& la,
                                        const uint32_t (&qf)[HD / 16][2]) {
  constexpr bool BF = (QT == 1 && KVF == 0);
  constexpr int KS = HD / 16;
  float s1[4] = {0.f, 0.f, 0.f, 0.f};
  s[0] = s[1] = s[2] = s[3] = 0.f;
  if constexpr (KVF == 0) {
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
      uint32_t a[4];
      ldmatrix_x4(a, kbase + swz(la.krow, 2 * ks + la.kch));
      if (ks & 1) mma_16816<BF>(s1, a, qf[ks][0], qf[ks][1]);
      else mma_16816<BF>(s, a, qf[ks][0], qf[ks][1]);
    }
  } else {
#pragma unroll
    for (int kp = 0; kp < KS / 2; ++kp) {
      uint32_t r[4];
      ldmatrix_x4(r, kbase + swz(la.krow, 2 * kp + la.kch));
      uint32_t a[4], b[4];
      a[0] = fp8x2_to_f16x2<KVF == 2>((uint16_t)(r[0] & 0xffffu));
      a[1] = fp8x2_to_f16x2<KVF == 2>((uint16_t)(r[1] & 0xffffu));
      a[2] = fp8x2_to_f16x2<KVF == 2>((uint16_t)(r[0] >> 16));
      a[3] = fp8x2_to_f16x2<KVF == 2>((uint16_t)(r[1] >> 16));
      b[0] = fp8x2_to_f16x2<KVF == 2>((uint16_t)(r[2] & 0xffffu));
      b[1] = fp8x2_to_f16x2<KVF == 2>((uint16_t)(r[3] & 0xffffu));
      b[2] = fp8x2_to_f16x2<KVF == 2>((uint16_t)(r[2] >> 16));
      b[3] = fp8x2_to_f16x2<KVF == 2>((uint16_t)(r[3] >> 16));
      mma_16816<false>(s, a, qf[2 * kp][0], qf[2 * kp][1]);
      mma_16816<false>(s1, b, qf[2 * kp + 1][0], qf[2 * kp + 1][1]);
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) s[i] += s1[i];
}

// O^T(HD x 8 heads) += V^T . P^T for this warp's 16 tokens; (pb0, pb1) = B fragments of P^T
template <int QT, int KVF, int HD>
__device__ __forceinline__ void pv_tile(float (&o)[HD / 16][4], uint32_t vbase,
                                        const LaneAddr& la, uint32_t pb0, uint32_t pb1) {
  constexpr bool BF = (QT == 1 && KVF == 0);
  constexpr int MT = HD / 16;
  if constexpr (KVF == 0) {
#pragma unroll
    for (int mt = 0; mt < MT; ++mt) {
      uint32_t a[4];
      ldmatrix_x4_trans(a, vbase + swz(la.vrow, 2 * mt + la.vch));
      mma_16816<BF>(o[mt], a, pb0, pb1);
    }
  } else {
#pragma unroll
    for (int mp = 0; mp < MT / 2; ++mp) {
      uint32_t r[4];
      ldmatrix_x4_trans(r, vbase + swz(la.vrow, 2 * mp + la.vch));
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        // r[2h]: tokens 0-7, r[2h+1]: tokens 8-15 of chunk 2mp+h; bytes
        // [tok 2t:d 2g, d 2g+1 | tok 2t+1: d 2g, d 2g+1] -> group by d
        const uint32_t x = prmt(r[2 * h], 0u, 0x3120u);
        const uint32_t y = prmt(r[2 * h + 1], 0u, 0x3120u);
        uint32_t a[4];
        a[0] = fp8x2_to_f16x2<KVF == 2>((uint16_t)(x & 0xffffu));  // row g   <-> d = 2g
        a[1] = fp8x2_to_f16x2<KVF == 2>((uint16_t)(x >> 16));      // row g+8 <-> d = 2g+1
        a[2] = fp8x2_to_f16x2<KVF == 2>((uint16_t)(y & 0xffffu));
        a[3] = fp8x2_to_f16x2<KVF == 2>((uint16_t)(y >> 16));
        mma_16816<false>(o[2 * mp + h], a, pb0, pb1);
      }
    }
  }
}

// head-dim index of accumulator row (mt, hi) for lane group g
template <int KVF>
__device__ __forceinline__ int o_dim(int mt, int hi, int g) {
  if constexpr (KVF == 0) return 16 * mt + g + 8 * hi;
  else return 16 * mt + 2 * g + hi;
}

// ---- consumer: 4 warps, online softmax + MMA ---------------------------------
template <int QT, int KVF, int HD, int LOADER>
__device__ __forceinline__ void consume(const DecodeParams& p, const SegSource& src,
                                        const Segment* seg_cache, uint32_t ring,
                                        uint32_t full_bar, uint32_t empty_bar, float* scratch_o,
                                        float* scratch_ml) {
  using G = TileGeom<KVF, HD>;
  constexpr bool BF = (QT == 1 && KVF == 0);  // operand type of P
  constexpr int MT = HD / 16;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t = lane & 3;
  const LaneAddr la = lane_addr<KVF>(warp, lane);
  const int GS = p.group_size;
  const bool use_cap = p.soft_cap > 0.f;
  const float cap_pre = use_cap ? p.sm_scale / p.soft_cap : 0.f;
  const float cap_post = use_cap ? p.soft_cap * 1.4426950408889634f : 0.f;

  int stage = 0;
  uint32_t phase = 0;
  for (int k = 0; k < src.count; ++k) {
    const Segment sg = seg_get(src, seg_cache, k);
    uint32_t qf[MT][2];
    {
      const char* qrow = static_cast<const char*>(p.q) +
                         ((int64_t)sg.request * p.q_stride_n +
                          (int64_t)(sg.kv_head * GS + g) * p.q_stride_h) * 2;
      load_q_frags<QT, KVF, HD>(qf, qrow, g < GS, t);
    }
    float o[MT][4];
#pragma unroll
    for (int i = 0; i < MT; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
    float m[2] = {kNegInf, kNegInf}, l[2] = {0.f, 0.f};

    for (int tile = sg.tile_begin; tile < sg.tile_end; ++tile) {
      mbar_wait(full_bar + 8 * stage, phase);
      const int tok0 = tile * kTileTokens + warp * 16;
      const int vcount = min(16, sg.tok_hi - tok0);
      if (vcount > 0 && tok0 + 16 > sg.tok_lo) {  // warp-uniform
        const uint32_t kbase = ring + stage * G::STAGE_BYTES;
        const uint32_t vbase = kbase + G::KV_BYTES;
        if constexpr (LOADER == 0) {
          if (vcount < 16) {  // rows the TMA did not (validly) fill: zero V so 0 * junk != NaN
            for (int i = lane; i < (16 - vcount) * G::CHUNKS; i += 32) {
              const int row = warp * 16 + vcount + i / G::CHUNKS;
              const uint32_t a = vbase + swz(row, i % G::CHUNKS);
              asm volatile("st.shared.v4.b32 [%0], {%1,%1,%1,%1};" ::"r"(a), "r"(0u) : "memory");
            }
            __syncwarp();
          }
        }
        float s[4];
        qk_tile<QT, KVF, HD>(s, kbase, la, qf);
        // s[0],s[1]: token g, heads 2t,2t+1;  s[2],s[3]: token g+8
        const int ta = tok0 + g, tb = ta + 8;
        const bool va = ta >= sg.tok_lo && ta < sg.tok_hi;
        const bool vb = tb >= sg.tok_lo && tb < sg.tok_hi;
        float x[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          x[i] = use_cap ? cap_post * tanhf_approx(s[i] * cap_pre) : s[i] * p.sm_scale_log2;
        }
        x[0] = va ? x[0] : kNegInf; x[1] = va ? x[1] : kNegInf;
        x[2] = vb ? x[2] : kNegInf; x[3] = vb ? x[3] : kNegInf;
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
        pr[0] = va ? ex2f(x[0] - m[0]) : 0.f;
        pr[1] = va ? ex2f(x[1] - m[1]) : 0.f;
        pr[2] = vb ? ex2f(x[2] - m[0]) : 0.f;
        pr[3] = vb ? ex2f(x[3] - m[1]) : 0.f;
        l[0] = l[0] * sc[0] + pr[0] + pr[2];
        l[1] = l[1] * sc[1] + pr[1] + pr[3];
#pragma unroll
        for (int i = 0; i < MT; ++i) {
          o[i][0] *= sc[0]; o[i][1] *= sc[1]; o[i][2] *= sc[0]; o[i][3] *= sc[1];
        }
        const uint32_t pb0 = movmatrix_trans(pack2<BF>(pr[0], pr[1]));
        const uint32_t pb1 = movmatrix_trans(pack2<BF>(pr[2], pr[3]));
        pv_tile<QT, KVF, HD>(o, vbase, la, pb0, pb1);
        if constexpr (LOADER == 0) {
          if (vcount < 16) fence_proxy_async_smem();  // generic zero-fill before next TMA write
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(empty_bar + 8 * stage);
      if (++stage == p.num_stages) { stage = 0; phase ^= 1; }
    }

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
    named_bar_sync(1, kConsumerWarps * 32);
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
        inv[j] = any ? (direct ? p.v_scale : 1.f) / L[j] : 0.f;
        lse2[j] = any ? M[j] + lg2f(L[j]) : kNegInf;
      }
      const int h0 = 2 * t;
      if (direct) {
        const int64_t hq = (int64_t)sg.request * p.num_qo_heads + sg.kv_head * GS;
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          if (h0 + j < GS) {
            if (QT == 1) {
              __nv_bfloat16* orow = static_cast<__nv_bfloat16*>(p.o) + (hq + h0 + j) * HD;
#pragma unroll
              for (int i = 0; i < MT; ++i) {
                orow[o_dim<KVF>(i, 0, g)] = __float2bfloat16_rn(o[i][j] * inv[j]);
                orow[o_dim<KVF>(i, 1, g)] = __float2bfloat16_rn(o[i][2 + j] * inv[j]);
              }
            } else {
              __half* orow = static_cast<__half*>(p.o) + (hq + h0 + j) * HD;
#pragma unroll
              for (int i = 0; i < MT; ++i) {
                orow[o_dim<KVF>(i, 0, g)] = __float2half_rn(o[i][j] * inv[j]);
                orow[o_dim<KVF>(i, 1, g)] = __float2half_rn(o[i][2 + j] * inv[j]);
              }
            }
            if (p.lse != nullptr && g == 0) p.lse[hq + h0 + j] = lse2[j];
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
      }
    }
  }
}

// threads per CTA: 4 consumer warps + 1 (TMA) or 4 (cp.async) producer warps
template <int LOADER>
struct BlockDim {
  static constexpr int kProducerWarps = (LOADER == 0) ? 1 : 4;
  static constexpr int kThreads = (kConsumerWarps + kProducerWarps) * 32;
};

// QT: 0 f16, 1 bf16 (q / o dtype).  KVF: 0 = KV has q's dtype, 1 = e4m3, 2 = e5m2.
// LOADER: 0 = TMA boxes, 1 = cp.async gathers.
template <int QT, int KVF, int HD, int LOADER>
__global__ void __launch_bounds__(BlockDim<LOADER>::kThreads)
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
  const SegSource src = seg_source_init(p);
  for (int i = tid; i < min(src.count, kSegCache); i += blockDim.x) seg_cache[i] = seg_make(src, i);
  __syncthreads();
  if (src.count == 0) return;
  if (p.enable_pdl) pdl_wait();  // q (and the KV cache) come from upstream kernels

  if (tid < kConsumerWarps * 32) {
    consume<QT, KVF, HD, LOADER>(p, src, seg_cache, ring, full_bar, empty_bar, scratch_o,
                                 scratch_ml);
  } else if (LOADER == 0) {
    produce_tma<KVF, HD>(p, src, seg_cache, &kmap, &vmap, ring, full_bar, empty_bar);
  } else {
    produce_cpasync<KVF, HD>(p, src, seg_cache, tid - kConsumerWarps * 32, ring, full_bar,
                             empty_bar);
  }
  if (p.enable_pdl) pdl_launch_dependents();
}

}  // namespace fib200
