// decode_kernel_tc5.cuh — fp8-KV paged decode on the 5th-generation tensor cores (tcgen05 + TMEM).
//
// sm_100a has no native fp8 mma.sync (ptxas lowers it to F2FP conversions + HMMA), so the
// register-level path pays one conversion per KV element and stalls near 5.3 TB/s.  Here the fp8
// K / V bytes that TMA wrote to shared memory (128B swizzle) are consumed as they are by
// tcgen05.mma.kind::f8f6f4 — zero per-element ALU work:
//
//   S[128 tokens x 32]  = K_tile[128 x 128 fp8, K-major SW128]  x  Qs^T   (4 MMAs, k = 32 each)
//       Qs = the 16-bit query split into three e4m3 terms (rows 8t+h: term t of head h) so that
//       q.k keeps ~12 significant bits; the three partial logits are summed after tcgen05.ld.
//   O[128 d x 16]      += V_tile^T[128 d x 128 tokens, MN-major SW128]  x  P[128 tokens x 16]
//       P = probabilities of the 8 heads as e4m3 hi (cols 0-7) + residual lo (cols 8-15), written
//       by the softmax threads as one 16-byte row per token (MN-major, no swizzle).
//
// Roles (192 threads): warps 0-3 = softmax (thread t <-> token t of the tile, TMEM lane t),
// warp 4 = TMA producer (same per-page boxes as the mma.sync kernel), warp 5 = MMA issuer (one
// elected lane).  S is double-buffered in TMEM, P in shared memory; QK of tile i+1 is issued
// before PV of tile i so the tensor pipe runs under the softmax of tile i.
// The softmax keeps ONE reference exponent rho[h] per head for the whole CTA (the accumulator
// in TMEM is shared): per tile the 4 warp maxima are exchanged through shared memory and rho is
// raised only when the tile maximum exceeds it by more than 2^5; then O is rescaled in TMEM
// (tcgen05.ld / st).  p = 2^(x - rho + 3) <= 256 fits e4m3; hi + lo keeps ~8 bits.
// Semantics (base-2 lse, -5e4 sentinel, empty -> o = 0) as in decode_kernel.cuh.
//
// HP (head-pair plans, DecodeParams::head_pairs): a unit is a request and a PAIR of adjacent KV heads,
// a tile 64 tokens of both.  One TMA box brings 8-row groups ordered (token / 8, head), so the 128 rows
// of a stage's K (and V) part alternate between the heads every 8 rows and softmax thread t owns
// token 8 (t / 16) + t % 8 of head (t / 8) % 2:
//   S[128 rows x 64]  = K rows x [Qs(head 0) | Qs(head 1)]^T     (one MMA chain; a row reads only the
//                        32 columns of its own head, the other 32 are never looked at)
//   O_h[128 d x 16]  += V rows of head h (8-row groups at twice the stride) x P rows of head h
// Everything per head — rho, l, the bump vote's maxima — is exchanged among the 64 threads of that head.
// The query buffer is single (8 KiB); the issuer hands it back through `q_free`.
#pragma once
#include "decode_kernel.cuh"
#include "tc5.cuh"

namespace fib200 {

constexpr int kTc5Threads = 192;
// TMEM columns: S[b] at kTc5ColS + sstride b (2 buffers), O after them
template <bool HP>  // true: 64 logit columns per row (head pairs, wide units)
struct Tc5Cols {
  static constexpr uint32_t kSStride = HP ? 64 : 32;
  static constexpr uint32_t kO = 2 * kSStride;          // HP: O_h at kO + 16 h
  static constexpr uint32_t kAlloc = HP ? 256 : 128;    // a power of two >= kO + (HP ? 32 : 16)
};
// kernel modes kTc5Single / kTc5Pairs / kTc5Wide: decode_params.h
constexpr int kTc5ColS = 0;
constexpr float kTc5Bump = 5.f;     // raise rho when a tile max exceeds it by more than 2^5
constexpr float kTc5PShift = 3.f;   // p = 2^(x - rho + 3): <= 2^8 right after a bump

// barrier slots after the ring's full/empty arrays (SmemLayout::BAR_BYTES reserves 128 bytes)
struct Tc5Bars {
  uint32_t s_ready[2], s_free[2], p_ready[2], p_free[2], q_ready[2], q_free;
};
__device__ __forceinline__ Tc5Bars tc5_bars(uint32_t base) {
  Tc5Bars b;
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    b.s_ready[i] = base + 8 * i;
    b.s_free[i] = base + 16 + 8 * i;
    b.p_ready[i] = base + 32 + 8 * i;
    b.p_free[i] = base + 48 + 8 * i;
    b.q_ready[i] = base + 64 + 8 * i;
  }
  b.q_free = base + 80;
  return b;
}

// ---- query preparation: 128 softmax threads write Qs (32 rows x 128 B, K-major SW128) ---------
// thread t: head h = t / 16, dims d0 = 8 (t % 16) .. d0 + 7.  Row 8 term + h holds term `term` of
// head h; rows 24-31 and heads >= group_size are zero.  dq[h] = amax / 256 is the inverse scale.
// HP: two passes, rows 32 hp + .. and dq[8 hp + h] for head hp of the pair.
template <int QT, bool HP>
__device__ __forceinline__ void tc5_prepare_q(const DecodeParams& p, const Segment& sg, uint8_t* qbuf,
                                              float* dq_out, int tid) {
  const int h = tid >> 4, d0 = (tid & 15) * 8;
  const int GS = p.group_size;                // query heads of the unit (both heads of a pair)
  const int G1 = HP ? GS >> 1 : GS;           // query heads per KV head
#pragma unroll
  for (int hp = 0; hp < (HP ? 2 : 1); ++hp) {
  float x[8];
  float amax = 0.f;
  if (h < G1) {
    const char* qrow = static_cast<const char*>(p.q) +
                       ((int64_t)sg.request * p.q_stride_n +
                        (int64_t)(sg.kv_head * GS + hp * G1 + h) * p.q_stride_h) * 2 + d0 * 2;
    const uint2 v0 = __ldg(reinterpret_cast<const uint2*>(qrow));
    const uint2 v1 = __ldg(reinterpret_cast<const uint2*>(qrow) + 1);
    const uint32_t w[4] = {v0.x, v0.y, v1.x, v1.y};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if constexpr (QT == 1) {
        x[2 * i] = __uint_as_float(w[i] << 16);
        x[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
      } else {
        const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
        x[2 * i] = f.x;
        x[2 * i + 1] = f.y;
      }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) amax = fmaxf(amax, fabsf(x[i]));
  } else {
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = 0.f;
  }
#pragma unroll
  for (int s = 1; s < 16; s <<= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, s));
  const bool nz = amax > 0.f && amax < 3.0e38f;
  const float sq = nz ? 256.f / amax : 1.f;
  if ((tid & 15) == 0) dq_out[8 * hp + h] = nz ? amax * (1.f / 256.f) : 1.f;
  uint32_t t[4][3];
#pragma unroll
  for (int i = 0; i < 4; ++i) split3_e4m3x2(t[i], x[2 * i] * sq, x[2 * i + 1] * sq);
  const int chunk = d0 >> 4, within = d0 & 15;
#pragma unroll
  for (int term = 0; term < 4; ++term) {
    const int row = 32 * hp + 8 * term + h;
    uint2 v = make_uint2(0u, 0u);
    if (term < 3) {
      v.x = t[0][term] | (t[1][term] << 16);
      v.y = t[2][term] | (t[3][term] << 16);
    }
    *reinterpret_cast<uint2*>(qbuf + swz<128>(row, chunk) + within) = v;
  }
  }
}

// head_dim 256: thread t: head h = t / 16, dims d0 = 16 (t % 16) .. d0 + 15 = one 16-byte chunk of fp8 terms.
// Qs is two 128-byte panels of 32 rows each (dims 0-127 | 128-255, 4 KiB apart).
template <int QT>
__device__ __forceinline__ void tc5_prepare_q256(const DecodeParams& p, const Segment& sg, uint8_t* qbuf,
                                                 float* dq_out, int tid) {
  const int h = tid >> 4, d0 = (tid & 15) * 16;
  const int GS = p.group_size;
  float x[16];
  float amax = 0.f;
  if (h < GS) {
    const char* qrow = static_cast<const char*>(p.q) +
                       ((int64_t)sg.request * p.q_stride_n +
                        (int64_t)(sg.kv_head * GS + h) * p.q_stride_h) * 2 + d0 * 2;
#pragma unroll
    for (int v = 0; v < 4; ++v) {
      const uint2 w2 = __ldg(reinterpret_cast<const uint2*>(qrow) + v);
      const uint32_t w[2] = {w2.x, w2.y};
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        if constexpr (QT == 1) {
          x[4 * v + 2 * i] = __uint_as_float(w[i] << 16);
          x[4 * v + 2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
        } else {
          const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
          x[4 * v + 2 * i] = f.x;
          x[4 * v + 2 * i + 1] = f.y;
        }
      }
    }
#pragma unroll
    for (int i = 0; i < 16; ++i) amax = fmaxf(amax, fabsf(x[i]));
  } else {
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = 0.f;
  }
#pragma unroll
  for (int s2 = 1; s2 < 16; s2 <<= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, s2));
  const bool nz = amax > 0.f && amax < 3.0e38f;
  const float sq = nz ? 256.f / amax : 1.f;
  if ((tid & 15) == 0) dq_out[h] = nz ? amax * (1.f / 256.f) : 1.f;
  uint32_t t[8][3];
#pragma unroll
  for (int i = 0; i < 8; ++i) split3_e4m3x2(t[i], x[2 * i] * sq, x[2 * i + 1] * sq);
  const int chunk = tid & 15;
#pragma unroll
  for (int term = 0; term < 4; ++term) {
    const int row = 8 * term + h;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (term < 3) {
      v.x = t[0][term] | (t[1][term] << 16);
      v.y = t[2][term] | (t[3][term] << 16);
      v.z = t[4][term] | (t[5][term] << 16);
      v.w = t[6][term] | (t[7][term] << 16);
    }
    *reinterpret_cast<uint4*>(qbuf + swz<32>(row, chunk)) = v;
  }
}

// ---- MMA issuer (warp 5; one elected lane issues, the warp stays converged) ---------------------
template <int KVF, int MODE, int HD = 128>
__device__ __forceinline__ void tc5_issue(const DecodeParams& p, const SegSource& src,
                                          uint32_t ring, uint32_t qbufs,
                                          uint32_t pbufs, uint32_t full_bar, uint32_t empty_bar,
                                          const Tc5Bars& bars, uint32_t tmem) {
  using G = TileGeom<KVF, HD>;
  constexpr bool HP = MODE == kTc5Pairs, WIDE = MODE == kTc5Wide;
  constexpr bool Q1 = MODE != kTc5Single;  // 64-row query buffer, single, handed back through q_free
  static_assert(MODE == kTc5Single || HD == 128, "head pairs / wide units: head_dim 128");
  constexpr int AF = (KVF == 2) ? 1 : 0;  // e5m2 : e4m3 storage
  using TC = Tc5Cols<Q1>;
  constexpr uint32_t idesc_qk = tc5::instr_desc_f8(AF, 0, /*a K-major*/ 0, /*b K-major*/ 0, 128, Q1 ? 64 : 32);
  constexpr uint32_t idesc_pv = tc5::instr_desc_f8(AF, 0, /*a MN-major*/ 1, /*b MN-major*/ 1, 128, WIDE ? 32 : 16);
  const bool leader = (threadIdx.x & 31) == 0;
  int stage = 0;
  uint32_t phase = 0;
  uint32_t n = 0;  // tiles issued so far by this CTA: S / P buffer = n & 1
  // pending P.V of the previous tile
  bool pend = false;
  int pend_stage = 0;
  uint32_t pend_n = 0, pend_acc = 0;

  auto issue_pv = [&]() {
    // WIDE: ONE probability buffer (4 KiB: 128 rows x [columns 0-7 hi lo | columns 8-15 hi lo]), phase = tile
    const uint32_t b = WIDE ? 0u : pend_n & 1u;
    mbar_wait(bars.p_ready[b], WIDE ? pend_n & 1u : (pend_n >> 1) & 1u);
    tc5::fence_after_sync();
    if (leader) {
      const uint32_t vbase = ring + pend_stage * G::STAGE_BYTES + G::KV_BYTES;
      const uint32_t pb = pbufs + b * 2048u;
      if constexpr (WIDE) {
        // N = 32: the two 16-byte column groups of a row are 2 KiB apart (SBO), 8-token groups 128 B (LBO)
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          const uint64_t ad = tc5::smem_desc(vbase + ks * 4096u, 1024u, 1024u, tc5::kLayoutSw128);
          const uint64_t bd = tc5::smem_desc(pb + ks * 512u, 128u, 2048u, tc5::kLayoutNone);
          tc5::mma_f8(tmem + TC::kO, ad, bd, idesc_pv, (ks > 0) ? 1u : pend_acc);
        }
      } else if constexpr (HP) {
        // head hp owns every other 8-row group: V groups 2 KiB apart, P groups 256 B apart
#pragma unroll
        for (int hp = 0; hp < 2; ++hp) {
#pragma unroll
          for (int ks = 0; ks < 2; ++ks) {  // 32 tokens of head hp per MMA
            const uint64_t ad = tc5::smem_desc(vbase + hp * 1024u + ks * 8192u, 1024u, 2048u, tc5::kLayoutSw128);
            const uint64_t bd = tc5::smem_desc(pb + hp * 128u + ks * 1024u, 256u, 256u, tc5::kLayoutNone);
            tc5::mma_f8(tmem + TC::kO + 16u * hp, ad, bd, idesc_pv, (ks > 0) ? 1u : pend_acc);
          }
        }
      } else {
#pragma unroll
        for (int mh = 0; mh < HD / 128; ++mh) {  // 128 output dims per accumulator (one V panel)
#pragma unroll
          for (int ks = 0; ks < 4; ++ks) {  // 32 tokens per MMA
            const uint64_t ad = tc5::smem_desc(vbase + mh * G::PANEL_BYTES + ks * 4096u, 1024u, 1024u,
                                               tc5::kLayoutSw128);
            const uint64_t bd = tc5::smem_desc(pb + ks * 512u, 128u, 128u, tc5::kLayoutNone);
            tc5::mma_f8(tmem + TC::kO + 16u * mh, ad, bd, idesc_pv, (ks > 0) ? 1u : pend_acc);
          }
        }
      }
      tc5::commit(empty_bar + 8 * pend_stage);  // K / V stage back to the TMA producer
      tc5::commit(bars.p_free[b]);              // P buffer reusable, O updated
    }
    __syncwarp();
    pend = false;
  };

  for (int k = 0;; ++k) {
    Segment sg;
    if (!seg_fetch(src, k, sg)) break;
    __syncwarp();
    if (leader) src.ctl[kQDone1] = k + 1;  // this warp holds its copy of segment k
    // every segment: never lag a phase behind
    if constexpr (Q1) mbar_wait(bars.q_ready[0], k & 1u);
    else mbar_wait(bars.q_ready[k & 1], (k >> 1) & 1u);
    const uint32_t qb = Q1 ? qbufs : qbufs + (k & 1) * (32u * HD);
    if (Q1 && leader && sg.tile_end <= sg.tile_begin) mbar_arrive(bars.q_free);  // nothing reads this query
    for (int tile = sg.tile_begin; tile < sg.tile_end; ++tile) {
      const uint32_t b = n & 1u;
      // wait for the next K tile (and a free S buffer), but serve the pending P.V the moment its
      // probabilities arrive: that frees the K / V stage for the TMA producer as early as possible
      {
        const long long t0 = clock64();
        while (true) {
          if (pend && mbar_try_wait(bars.p_ready[WIDE ? 0u : pend_n & 1u], WIDE ? pend_n & 1u : (pend_n >> 1) & 1u))
            issue_pv();
          if (mbar_try_wait(full_bar + 8 * stage, phase) &&
              mbar_try_wait(bars.s_free[b], ((n >> 1) & 1u) ^ 1u))
            break;
          if (clock64() - t0 > 4000000000LL) {
            printf("fi_b200: tcgen05 issuer timed out (block %d tile %u)\n", blockIdx.x, n);
            __trap();
          }
        }
      }
      tc5::fence_after_sync();
      if (leader) {
        const uint32_t kbase = ring + stage * G::STAGE_BYTES;
#pragma unroll
        for (int ks = 0; ks < HD / 32; ++ks) {  // 32 head dims per MMA: +32 bytes inside the SW128 row
          const uint32_t pn = ks >> 2, off = (ks & 3) * 32u;  // 128-dim panel (K: PANEL_BYTES, Qs: 4 KiB apart)
          const uint64_t ad = tc5::smem_desc(kbase + pn * G::PANEL_BYTES + off, 16u, 1024u, tc5::kLayoutSw128);
          const uint64_t bd = tc5::smem_desc(qb + pn * 4096u + off, 16u, 1024u, tc5::kLayoutSw128);
          tc5::mma_f8(tmem + kTc5ColS + TC::kSStride * b, ad, bd, idesc_qk, ks > 0 ? 1u : 0u);
        }
        tc5::commit(bars.s_ready[b]);
        if (Q1 && tile == sg.tile_end - 1) tc5::commit(bars.q_free);  // the single query buffer is free again
      }
      __syncwarp();
      if (pend) issue_pv();
      pend = true;
      pend_stage = stage;
      pend_n = n;
      pend_acc = (tile != sg.tile_begin) ? 1u : 0u;
      ++n;
      if (++stage == p.num_stages) { stage = 0; phase ^= 1; }
    }
    if (pend) issue_pv();  // the softmax warps need O complete before the segment epilogue
  }
}


// ---- softmax warps (0-3), one KV head per unit: thread tid <-> token tid of each tile, TMEM lane tid ----
// (kept apart from the head-pair version below: folding the two into one template cost the single-head
// kernel 1.7 % on C3 through instruction scheduling alone)
template <int QT, int KVF, int HD = 128>
__device__ __forceinline__ void tc5_softmax_single(const DecodeParams& p, const SegSource& src,
                                            uint8_t* smem_ring,
                                            uint8_t* qbufs, uint8_t* pbufs, float* scratch,
                                            const Tc5Bars& bars, uint32_t tmem, int num_stages) {
  using G = TileGeom<KVF, HD>;
  constexpr int NA = HD / 128;  // O accumulators of 128 dims each (thread tid: dims tid, 128 + tid)
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int GS = p.group_size;
  const bool use_cap = p.soft_cap > 0.f;
  const float cap_pre = use_cap ? p.sm_scale / p.soft_cap : 0.f;
  const float cap_post = use_cap ? p.soft_cap * 1.4426950408889634f : 0.f;
  const uint32_t lane_base = (uint32_t)(warp * 32) << 16;  // TMEM lane field
  // scratch (floats): wmax[2][4][8] @0, merge flag @64, dq[2][8] @72, lsum[4][8] @88
  float* wmax = scratch;
  volatile int* flag = reinterpret_cast<volatile int*>(scratch + 64);
  float* dqs = scratch + 72;
  float* lsum = scratch + 88;
  volatile int* vote = reinterpret_cast<volatile int*>(scratch + 124);  // [2][4]

  int stage = 0;
  uint32_t n = 0;  // tiles consumed so far: S / P buffer = n & 1
  PendingMerge pm;
  pm.mu = -1; pm.old = pm.cnt = pm.slot = pm.req = pm.head = 0;
  // The query of segment k + 1 is prepared during segment k when that segment is already
  // published (so the MMA warp can run ahead), else at the top of segment k + 1.  The decision
  // must be the same in all 128 threads: thread 0 peeks, the barrier broadcasts.
  bool have_q = false;
  for (int k = 0;; ++k) {
    Segment sg;
    if (!seg_fetch(src, k, sg)) break;
    if (!have_q) {
      if constexpr (HD == 256) tc5_prepare_q256<QT>(p, sg, qbufs + (k & 1) * 8192, dqs + (k & 1) * 8, tid);
      else tc5_prepare_q<QT, false>(p, sg, qbufs + (k & 1) * 4096, dqs + (k & 1) * 8, tid);
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(bars.q_ready[k & 1]);
    }
    if (tid == 0) {
      src.ctl[kQPeek] = seg_peek(src, k + 1) ? 1 : 0;
      src.ctl[kQDone0] = k;  // every thread is past segment k - 1 (barrier 3 of its epilogue)
    }
    named_bar_sync(1, 128);  // dq of segment k visible; scratch of segment k-1 no longer read
    float mul[8], pre[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float d = dqs[(k & 1) * 8 + j];
      mul[j] = d * p.sm_scale_log2;
      pre[j] = d * cap_pre;
    }
    float alc[8];  // ALiBi slope per head, in the units of the expression it is added to
#pragma unroll
    for (int j = 0; j < 8; ++j) alc[j] = 0.f;
    if (p.alibi_slopes != nullptr) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float a = j < GS ? __ldg(p.alibi_slopes + sg.kv_head * GS + j) : 0.f;
        alc[j] = a * (use_cap ? cap_pre : 1.4426950408889634f);
        pre[j] *= p.sm_scale;
      }
    }
    have_q = src.ctl[kQPeek] != 0;
    if (have_q) {
      const Segment sn = seg_read(src, k + 1);
      if constexpr (HD == 256) tc5_prepare_q256<QT>(p, sn, qbufs + ((k + 1) & 1) * 8192, dqs + ((k + 1) & 1) * 8, tid);
      else tc5_prepare_q<QT, false>(p, sn, qbufs + ((k + 1) & 1) * 4096, dqs + ((k + 1) & 1) * 8, tid);
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(bars.q_ready[(k + 1) & 1]);
    }
    float rho[8], l[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { rho[j] = kNegInf; l[j] = 0.f; }

    for (int tile = sg.tile_begin; tile < sg.tile_end; ++tile) {
      const uint32_t b = n & 1u;
      const int tok = tile * G::TT + tid;
      const bool valid = tok >= sg.tok_lo && tok < sg.tok_hi;
      float x[8];
      {
        mbar_wait(bars.s_ready[b], (n >> 1) & 1u);
        FI_TRACE_IF(tid == 0 && n == 0, 5);
        tc5::fence_after_sync();
        float c[32];
        tc5::ld_x32(tmem + lane_base + kTc5ColS + 32u * b, c);
        tc5::wait_ld();
        tc5::fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(bars.s_free[b]);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float sdot = c[j] + c[8 + j] + c[16 + j];
          const float v = use_cap ? cap_post * tanhf_approx(sdot * pre[j] + alc[j] * (float)tok)
                                  : sdot * mul[j] + alc[j] * (float)tok;
          x[j] = valid ? v : kNegInf;
        }
      }
      // Does any logit of this tile exceed rho by more than 2^kTc5Bump?  One vote per warp and one
      // CTA barrier in the common case; the exact per-head maxima are only reduced when rho moves.
      bool over = false;
#pragma unroll
      for (int j = 0; j < 8; ++j) over |= x[j] > rho[j] + kTc5Bump;
      const bool wover = __any_sync(0xffffffffu, over);
      if (lane == 0) vote[b * 4 + warp] = wover ? 1 : 0;
      named_bar_sync(2, 128);
      const bool bump = (vote[b * 4] | vote[b * 4 + 1] | vote[b * 4 + 2] | vote[b * 4 + 3]) != 0;
      float mt[8];
      if (bump) {  // CTA-uniform
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float m = x[j];
#pragma unroll
          for (int sft = 16; sft > 0; sft >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, sft));
          mt[j] = m;
        }
        float* wm = wmax + b * 32;
        if (lane < 8) {
          float mine = mt[0];
#pragma unroll
          for (int j = 1; j < 8; ++j) mine = lane == j ? mt[j] : mine;
          wm[warp * 8 + lane] = mine;
        }
        named_bar_sync(3, 128);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          mt[j] = fmaxf(fmaxf(wm[j], wm[8 + j]), fmaxf(wm[16 + j], wm[24 + j]));
      }
      if (bump) {  // CTA-uniform: every thread sees the same maxima
        float alpha[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float rn = mt[j] > rho[j] + kTc5Bump ? mt[j] : rho[j];
          alpha[j] = ex2f(rho[j] - rn);
          l[j] *= alpha[j];
          rho[j] = rn;
        }
        if (tile != sg.tile_begin) {
          // O in TMEM is relative to the old rho: wait for P.V of the previous tile, rescale
          mbar_wait(bars.p_free[b ^ 1u], ((n - 1) >> 1) & 1u);
          tc5::fence_after_sync();
#pragma unroll
          for (int a = 0; a < NA; ++a) {
            float o[16];
            tc5::ld_x16(tmem + lane_base + Tc5Cols<false>::kO + 16u * a, o);
            tc5::wait_ld();
#pragma unroll
            for (int j = 0; j < 8; ++j) { o[j] *= alpha[j]; o[8 + j] *= alpha[j]; }
            tc5::st_x16(tmem + lane_base + Tc5Cols<false>::kO + 16u * a, o);
          }
          tc5::wait_st();
          tc5::fence_before_sync();
        }
      }
      // probabilities -> e4m3 hi / lo, one 16-byte row per token
      uint32_t w[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float p0 = valid ? ex2f(x[2 * i] - rho[2 * i] + kTc5PShift) : 0.f;
        const float p1 = valid ? ex2f(x[2 * i + 1] - rho[2 * i + 1] + kTc5PShift) : 0.f;
        l[2 * i] += p0;
        l[2 * i + 1] += p1;
        w[i] = split_e4m3x2(p0, p1);  // hi pair in bits 0-15, lo pair in bits 16-31
      }
      uint4 row;
      row.x = (w[0] & 0xffffu) | (w[1] << 16);
      row.y = (w[2] & 0xffffu) | (w[3] << 16);
      row.z = (w[0] >> 16) | (w[1] & 0xffff0000u);
      row.w = (w[2] >> 16) | (w[3] & 0xffff0000u);
      mbar_wait(bars.p_free[b], ((n >> 1) & 1u) ^ 1u);  // P.V of tile n-2 done with this buffer
      *reinterpret_cast<uint4*>(pbufs + b * 2048 + tid * 16) = row;
      if (tok >= sg.tok_hi) {
        // this token's V row was not (validly) loaded: zero it so that 0 * junk cannot be NaN
#pragma unroll
        for (int pn = 0; pn < G::NP; ++pn) {
          uint4* vrow = reinterpret_cast<uint4*>(smem_ring + stage * G::STAGE_BYTES + G::KV_BYTES +
                                                 pn * G::PANEL_BYTES + tid * 128);
#pragma unroll
          for (int i = 0; i < 8; ++i) vrow[i] = make_uint4(0u, 0u, 0u, 0u);
        }
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(bars.p_ready[b]);
      ++n;
      if (++stage == num_stages) stage = 0;
    }

    FI_TRACE_IF(tid == 0, 6);
    // ---- segment epilogue: thread tid <-> head-dim index d = tid (TMEM lane of O) ---------------
    float o8[NA * 8];
    if (sg.tile_end > sg.tile_begin) {
      mbar_wait(bars.p_free[(n - 1) & 1u], ((n - 1) >> 1) & 1u);  // P.V of the last tile complete
      tc5::fence_after_sync();
#pragma unroll
      for (int a = 0; a < NA; ++a) {
        float o[16];
        tc5::ld_x16(tmem + lane_base + Tc5Cols<false>::kO + 16u * a, o);
        tc5::wait_ld();
#pragma unroll
        for (int j = 0; j < 8; ++j) o8[8 * a + j] = o[j] + o[8 + j];
      }
      tc5::fence_before_sync();
    } else {
#pragma unroll
      for (int j = 0; j < NA * 8; ++j) o8[j] = 0.f;
    }
    float L[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float s = l[j];
#pragma unroll
      for (int sft = 16; sft > 0; sft >>= 1) s += __shfl_xor_sync(0xffffffffu, s, sft);
      L[j] = s;
    }
    if (lane < 8) {
      float mine = L[0];
#pragma unroll
      for (int j = 1; j < 8; ++j) mine = lane == j ? L[j] : mine;
      lsum[warp * 8 + lane] = mine;
    }
    if (tid == 0) merge_post(p, pm, scratch);  // arrival of the previous partial
    named_bar_sync(3, 128);
    merge_run_posted<QT, HD>(p, scratch, tid);
#pragma unroll
    for (int j = 0; j < 8; ++j) L[j] = (lsum[j] + lsum[8 + j]) + (lsum[16 + j] + lsum[24 + j]);

    const bool direct = sg.slot < 0;
    const int64_t hq = (int64_t)sg.request * p.num_qo_heads + sg.kv_head * GS;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (j < GS) {
        const bool any = L[j] > 0.f;
        if (direct && any && p.sinks != nullptr)  // L is relative to rho and carries 2^kTc5PShift
          L[j] += ex2f(__ldg(p.sinks + sg.kv_head * GS + j) * 1.4426950408889634f - rho[j] + kTc5PShift);
        // numerator and denominator both carry the 2^kTc5PShift scale of p
        const float nrm = (direct ? p.v_scale : 1.f) / L[j];
        const float lse2 = any ? rho[j] + lg2f(L[j]) - kTc5PShift : kNegInf;
#pragma unroll
        for (int a = 0; a < NA; ++a) {
          const float val = any ? o8[8 * a + j] * nrm : 0.f;
          const int d = 128 * a + tid;
          if (direct) {
            if (QT == 1) static_cast<__nv_bfloat16*>(p.o)[(hq + j) * HD + d] = __float2bfloat16_rn(val);
            else static_cast<__half*>(p.o)[(hq + j) * HD + d] = __float2half_rn(val);
          } else {
            p.tmp_o[((int64_t)sg.slot * GS + j) * HD + d] = val;
          }
        }
        if (direct) {
          if (p.lse != nullptr && tid == j) p.lse[hq + j] = lse2;
        } else if (tid == j) {
          p.tmp_lse[(int64_t)sg.slot * GS + j] = lse2;
        }
      }
    }
    FI_TRACE_IF(tid == 0, 7);
    if (sg.merge_unit >= 0) {  // CTA-uniform: count the arrival, look at the result later
      named_bar_sync(2, 128);  // all 128 threads wrote their part of the partial
      if (tid == 0) merge_arrive(p, pm, sg);
    }
  }
  // end of this CTA's sequence: the last arrival is still unchecked
  if (tid == 0) merge_post(p, pm, scratch);
  named_bar_sync(3, 128);
  merge_run_posted<QT, HD>(p, scratch, tid);
  FI_TRACE_IF(tid == 0, 8);
}

// ---- softmax warps (0-3): thread tid <-> row tid of each tile (TMEM lane tid) ------------------
// row tid = token tid of the tile; HP: token 8 (tid / 16) + tid % 8 of head (tid / 8) % 2.
// reductions over the threads of one head: all 32 lanes; HP: the 16 lanes with the same bit 3
template <bool HP>
__device__ __forceinline__ float tc5_head_max(float m) {
  m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 16));
  if constexpr (!HP) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 8));
  m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 4));
  m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 2));
  return fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 1));
}
template <bool HP>
__device__ __forceinline__ float tc5_head_sum(float s) {
  s += __shfl_xor_sync(0xffffffffu, s, 16);
  if constexpr (!HP) s += __shfl_xor_sync(0xffffffffu, s, 8);
  s += __shfl_xor_sync(0xffffffffu, s, 4);
  s += __shfl_xor_sync(0xffffffffu, s, 2);
  return s + __shfl_xor_sync(0xffffffffu, s, 1);
}
// v[lane & 7] without a local-memory array index
__device__ __forceinline__ float tc5_pick8(const float (&v)[8], int lane) {
  float mine = v[0];
#pragma unroll
  for (int j = 1; j < 8; ++j) mine = (lane & 7) == j ? v[j] : mine;
  return mine;
}

template <int QT, int KVF, bool HP>
__device__ __forceinline__ void tc5_softmax(const DecodeParams& p, const SegSource& src,
                                            uint8_t* smem_ring,
                                            uint8_t* qbufs, uint8_t* pbufs, float* scratch,
                                            float* pair_scratch, const Tc5Bars& bars, uint32_t tmem,
                                            int num_stages) {
  using G = TileGeom<KVF, 128>;
  using TC = Tc5Cols<HP>;
  constexpr int HD = 128;
  constexpr int TT = HP ? G::TT / 2 : G::TT;  // tokens per tile
  constexpr int NH = HP ? 2 : 1;              // KV heads per unit
  constexpr int WS = 8 * NH;                  // per-warp exchange stride (floats)
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int GS = p.group_size;                // query heads of the unit
  const int G1 = HP ? GS >> 1 : GS;           // query heads per KV head
  const int hp = HP ? (tid >> 3) & 1 : 0;     // this row's head of the pair
  const int tok_in_tile = HP ? ((tid >> 4) << 3) | (tid & 7) : tid;
  const bool use_cap = p.soft_cap > 0.f;
  const float cap_pre = use_cap ? p.sm_scale / p.soft_cap : 0.f;
  const float cap_post = use_cap ? p.soft_cap * 1.4426950408889634f : 0.f;
  const uint32_t lane_base = (uint32_t)(warp * 32) << 16;  // TMEM lane field
  // scratch (floats): wmax @0 ([2][4][8]; HP: [4][16], single: a bump is followed by the next tile's
  // vote barrier before wmax is written again), merge flag @64, dq @72 ([2][8]; HP: [16]), lsum @88
  // ([4][8]), tmem slot @120, vote @124; HP, in pair_scratch: lsum [4][16] @0, alpha @64 [16], rho @80 [16]
  float* wmax = scratch;
  float* dqs = scratch + 72;
  float* lsum = HP ? pair_scratch : scratch + 88;
  float* alpha_sm = pair_scratch + 64;
  float* rho_sm = pair_scratch + 80;
  volatile int* vote = reinterpret_cast<volatile int*>(scratch + 124);  // [2][4]

  int stage = 0;
  uint32_t n = 0;  // tiles consumed so far: S / P buffer = n & 1
  PendingMerge pm;
  pm.mu = -1; pm.old = pm.cnt = pm.slot = pm.req = pm.head = 0;
  // The query of segment k + 1 is prepared during segment k when that segment is already
  // published (so the MMA warp can run ahead), else at the top of segment k + 1.  The decision
  // must be the same in all 128 threads: thread 0 peeks, the barrier broadcasts.
  // HP (one query buffer): "during" = after the last S tile of segment k has been read, once the
  // issuer's q_free commit says that every Q.K MMA of segment k is complete.
  bool have_q = false;
  for (int k = 0;; ++k) {
    Segment sg;
    if (!seg_fetch(src, k, sg)) break;
    if (!have_q) {
      if (HP && k > 0) mbar_wait(bars.q_free, (k - 1) & 1u);
      tc5_prepare_q<QT, HP>(p, sg, HP ? qbufs : qbufs + (k & 1) * 4096, HP ? dqs : dqs + (k & 1) * 8, tid);
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(bars.q_ready[HP ? 0 : k & 1]);
    }
    if (tid == 0) {
      src.ctl[kQPeek] = seg_peek(src, k + 1) ? 1 : 0;
      src.ctl[kQDone0] = k;  // every thread is past segment k - 1 (barrier 3 of its epilogue)
    }
    named_bar_sync(1, 128);  // dq of segment k visible; scratch of segment k-1 no longer read
    float mul[8], pre[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float d = dqs[(HP ? 8 * hp : (k & 1) * 8) + j];
      mul[j] = d * p.sm_scale_log2;
      pre[j] = d * cap_pre;
    }
    float alc[8];  // ALiBi slope per head, in the units of the expression it is added to
#pragma unroll
    for (int j = 0; j < 8; ++j) alc[j] = 0.f;
    if (p.alibi_slopes != nullptr) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float a = j < G1 ? __ldg(p.alibi_slopes + sg.kv_head * GS + hp * G1 + j) : 0.f;
        alc[j] = a * (use_cap ? cap_pre : 1.4426950408889634f);
        pre[j] *= p.sm_scale;
      }
    }
    have_q = src.ctl[kQPeek] != 0;
    if (!HP && have_q) {
      const Segment sn = seg_read(src, k + 1);
      tc5_prepare_q<QT, HP>(p, sn, qbufs + ((k + 1) & 1) * 4096, dqs + ((k + 1) & 1) * 8, tid);
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(bars.q_ready[(k + 1) & 1]);
    }
    float rho[8], l[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { rho[j] = kNegInf; l[j] = 0.f; }

    for (int tile = sg.tile_begin; tile < sg.tile_end; ++tile) {
      const uint32_t b = n & 1u;
      const int tok = tile * TT + tok_in_tile;
      const bool valid = tok >= sg.tok_lo && tok < sg.tok_hi;
      float x[8];
      {
        mbar_wait(bars.s_ready[b], (n >> 1) & 1u);
        FI_TRACE_IF(tid == 0 && n == 0, 5);
        tc5::fence_after_sync();
        // tcgen05.ld is warp-collective (one address per warp), and a warp holds rows of both heads
        // of a pair: both heads' 24 term columns are read and every row keeps its own
        float sd[8];
        if constexpr (HP) {
          float c0[16], c1[8], e0[16], e1[8];
          const uint32_t sb = tmem + lane_base + kTc5ColS + TC::kSStride * b;
          tc5::ld_x16(sb, c0);
          tc5::ld_x8(sb + 16u, c1);
          tc5::ld_x16(sb + 32u, e0);
          tc5::ld_x8(sb + 48u, e1);
          tc5::wait_ld();
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float s0 = c0[j] + c0[8 + j] + c1[j], s1 = e0[j] + e0[8 + j] + e1[j];
            sd[j] = hp ? s1 : s0;
          }
        } else {
          float c[32];
          tc5::ld_x32(tmem + lane_base + kTc5ColS + TC::kSStride * b, c);
          tc5::wait_ld();
#pragma unroll
          for (int j = 0; j < 8; ++j) sd[j] = c[j] + c[8 + j] + c[16 + j];
        }
        tc5::fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(bars.s_free[b]);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float sdot = sd[j];
          const float v = use_cap ? cap_post * tanhf_approx(sdot * pre[j] + alc[j] * (float)tok)
                                  : sdot * mul[j] + alc[j] * (float)tok;
          x[j] = valid ? v : kNegInf;
        }
      }
      // Does any logit of this tile exceed rho by more than 2^kTc5Bump?  One vote per warp and one
      // CTA barrier in the common case; the exact per-head maxima are only reduced when rho moves.
      bool over = false;
#pragma unroll
      for (int j = 0; j < 8; ++j) over |= x[j] > rho[j] + kTc5Bump;
      const bool wover = __any_sync(0xffffffffu, over);
      if (lane == 0) vote[b * 4 + warp] = wover ? 1 : 0;
      named_bar_sync(2, 128);
      const bool bump = (vote[b * 4] | vote[b * 4 + 1] | vote[b * 4 + 2] | vote[b * 4 + 3]) != 0;
      if (bump) {  // CTA-uniform: every thread of a head sees the same maxima
        float mt[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) mt[j] = tc5_head_max<HP>(x[j]);
        float* wm = HP ? wmax : wmax + b * 32;
        if (lane < WS) wm[warp * WS + lane] = tc5_pick8(mt, lane);  // HP: lanes 8-15 belong to head 1
        named_bar_sync(3, 128);
        float alpha[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int i = 8 * hp + j;
          const float m = fmaxf(fmaxf(wm[i], wm[WS + i]), fmaxf(wm[2 * WS + i], wm[3 * WS + i]));
          const float rn = m > rho[j] + kTc5Bump ? m : rho[j];
          alpha[j] = ex2f(rho[j] - rn);
          l[j] *= alpha[j];
          rho[j] = rn;
        }
        if (tile != sg.tile_begin) {
          // O in TMEM is relative to the old rho: wait for P.V of the previous tile, rescale.  This
          // thread is row d of O^T for ALL heads of the unit.
          if constexpr (HP) {
            if (tid < 16) alpha_sm[tid] = tc5_pick8(alpha, tid);  // tid / 8 = that thread's head
            named_bar_sync(3, 128);
          }
          mbar_wait(bars.p_free[b ^ 1u], ((n - 1) >> 1) & 1u);
          tc5::fence_after_sync();
#pragma unroll
          for (int h2 = 0; h2 < NH; ++h2) {
            float o[16];
            tc5::ld_x16(tmem + lane_base + TC::kO + 16u * h2, o);
            tc5::wait_ld();
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float a = HP ? alpha_sm[8 * h2 + j] : alpha[j];
              o[j] *= a;
              o[8 + j] *= a;
            }
            tc5::st_x16(tmem + lane_base + TC::kO + 16u * h2, o);
          }
          tc5::wait_st();
          tc5::fence_before_sync();
        }
      }
      // probabilities -> e4m3 hi / lo, one 16-byte row per token
      uint32_t w[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float p0 = valid ? ex2f(x[2 * i] - rho[2 * i] + kTc5PShift) : 0.f;
        const float p1 = valid ? ex2f(x[2 * i + 1] - rho[2 * i + 1] + kTc5PShift) : 0.f;
        l[2 * i] += p0;
        l[2 * i + 1] += p1;
        w[i] = split_e4m3x2(p0, p1);  // hi pair in bits 0-15, lo pair in bits 16-31
      }
      uint4 row;
      row.x = (w[0] & 0xffffu) | (w[1] << 16);
      row.y = (w[2] & 0xffffu) | (w[3] << 16);
      row.z = (w[0] >> 16) | (w[1] & 0xffff0000u);
      row.w = (w[2] >> 16) | (w[3] & 0xffff0000u);
      mbar_wait(bars.p_free[b], ((n >> 1) & 1u) ^ 1u);  // P.V of tile n-2 done with this buffer
      *reinterpret_cast<uint4*>(pbufs + b * 2048 + tid * 16) = row;
      if (tok >= sg.tok_hi) {
        // this token's V row was not (validly) loaded: zero it so that 0 * junk cannot be NaN
        uint4* vrow = reinterpret_cast<uint4*>(smem_ring + stage * G::STAGE_BYTES + G::KV_BYTES +
                                               tid * 128);
#pragma unroll
        for (int i = 0; i < 8; ++i) vrow[i] = make_uint4(0u, 0u, 0u, 0u);
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(bars.p_ready[b]);
      ++n;
      if (++stage == num_stages) stage = 0;
    }
    if (HP && have_q) {  // CTA-uniform; see the comment at have_q
      // a segment without tiles has no barrier since dq of segment k was read: it is overwritten below
      if (sg.tile_end <= sg.tile_begin) named_bar_sync(2, 128);
      mbar_wait(bars.q_free, k & 1u);
      const Segment sn = seg_read(src, k + 1);
      tc5_prepare_q<QT, HP>(p, sn, qbufs, dqs, tid);
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(bars.q_ready[0]);
    }

    FI_TRACE_IF(tid == 0, 6);
    // ---- segment epilogue: thread tid <-> head-dim index d = tid (TMEM lane of O) ---------------
    float o8[NH * 8];
    if (sg.tile_end > sg.tile_begin) {
      mbar_wait(bars.p_free[(n - 1) & 1u], ((n - 1) >> 1) & 1u);  // P.V of the last tile complete
      tc5::fence_after_sync();
#pragma unroll
      for (int h2 = 0; h2 < NH; ++h2) {
        float o[16];
        tc5::ld_x16(tmem + lane_base + TC::kO + 16u * h2, o);
        tc5::wait_ld();
#pragma unroll
        for (int j = 0; j < 8; ++j) o8[8 * h2 + j] = o[j] + o[8 + j];
      }
      tc5::fence_before_sync();
    } else {
#pragma unroll
      for (int i = 0; i < NH * 8; ++i) o8[i] = 0.f;
    }
    {
      float L[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) L[j] = tc5_head_sum<HP>(l[j]);
      if (lane < WS) lsum[warp * WS + lane] = tc5_pick8(L, lane);
      if (HP && tid < 16) rho_sm[tid] = tc5_pick8(rho, tid);
    }
    if (tid == 0) merge_post(p, pm, scratch);  // arrival of the previous partial
    named_bar_sync(3, 128);
    merge_run_posted<QT, HD>(p, scratch, tid);

    const bool direct = sg.slot < 0;
    const int64_t hq = (int64_t)sg.request * p.num_qo_heads + sg.kv_head * GS;
#pragma unroll
    for (int i = 0; i < NH * 8; ++i) {  // i: exchange index = 8 * (head of the pair) + head in its group
      {
        const int h2 = i >> 3, j = i & 7;
        if (j < G1) {
          const int qh = h2 * G1 + j;      // query head inside the unit
          float Lj = (lsum[i] + lsum[WS + i]) + (lsum[2 * WS + i] + lsum[3 * WS + i]);
          const float rj = HP ? rho_sm[i] : rho[j];
          const bool any = Lj > 0.f;
          if (direct && any && p.sinks != nullptr)  // L is relative to rho and carries 2^kTc5PShift
            Lj += ex2f(__ldg(p.sinks + sg.kv_head * GS + qh) * 1.4426950408889634f - rj + kTc5PShift);
          // numerator and denominator both carry the 2^kTc5PShift scale of p
          const float val = any ? o8[i] * ((direct ? p.v_scale : 1.f) / Lj) : 0.f;
          const float lse2 = any ? rj + lg2f(Lj) - kTc5PShift : kNegInf;
          if (direct) {
            if (QT == 1) static_cast<__nv_bfloat16*>(p.o)[(hq + qh) * HD + tid] = __float2bfloat16_rn(val);
            else static_cast<__half*>(p.o)[(hq + qh) * HD + tid] = __float2half_rn(val);
            if (p.lse != nullptr && tid == i) p.lse[hq + qh] = lse2;
          } else {
            p.tmp_o[((int64_t)sg.slot * GS + qh) * HD + tid] = val;
            if (tid == i) p.tmp_lse[(int64_t)sg.slot * GS + qh] = lse2;
          }
        }
      }
    }
    FI_TRACE_IF(tid == 0, 7);
    if (sg.merge_unit >= 0) {  // CTA-uniform: count the arrival, look at the result later
      named_bar_sync(2, 128);  // all 128 threads wrote their part of the partial
      if (tid == 0) merge_arrive(p, pm, sg);
    }
  }
  // end of this CTA's sequence: the last arrival is still unchecked
  if (tid == 0) merge_post(p, pm, scratch);
  named_bar_sync(3, 128);
  merge_run_posted<QT, HD>(p, scratch, tid);
  FI_TRACE_IF(tid == 0, 8);
}

// ---- wide units (kTc5Wide): ONE KV head, up to 16 (row, head) query columns per unit ------------
// Columns map to (row, head) through row_head (decode_kernel.cuh); the query buffer holds two
// 32-row groups (columns 0-7, 8-15), S has 64 columns per token, and a softmax thread — still one
// token — works through the two column groups one after the other (rho / l for 16 columns stay in
// registers, the per-column scales and causal cuts come from shared memory).  P is ONE 4 KiB buffer:
// the thread waits for P.V of the previous tile before it overwrites its row.
template <int QT>
__device__ __forceinline__ void tc5_prepare_q_wide(const DecodeParams& p, const Segment& sg, uint8_t* qbuf,
                                                   float* dq_out, int tid) {
  const int h = tid >> 4, d0 = (tid & 15) * 8;
  const int GS = p.group_size;
#pragma unroll
  for (int cg = 0; cg < 2; ++cg) {
    const int col = 8 * cg + h;
    float x[8];
    float amax = 0.f;
    if (col < GS) {
      const RowHead rh = row_head(p, sg.kv_head, col);
      const char* qrow = static_cast<const char*>(p.q) +
                         (rh.token(p, sg.request) * p.q_stride_n + (int64_t)rh.head * p.q_stride_h) * 2 + d0 * 2;
      const uint2 v0 = __ldg(reinterpret_cast<const uint2*>(qrow));
      const uint2 v1 = __ldg(reinterpret_cast<const uint2*>(qrow) + 1);
      const uint32_t w[4] = {v0.x, v0.y, v1.x, v1.y};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        if constexpr (QT == 1) {
          x[2 * i] = __uint_as_float(w[i] << 16);
          x[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
        } else {
          const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
          x[2 * i] = f.x;
          x[2 * i + 1] = f.y;
        }
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) amax = fmaxf(amax, fabsf(x[i]));
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) x[i] = 0.f;
    }
#pragma unroll
    for (int s2 = 1; s2 < 16; s2 <<= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, s2));
    const bool nz = amax > 0.f && amax < 3.0e38f;
    const float sq = nz ? 256.f / amax : 1.f;
    if ((tid & 15) == 0) dq_out[col] = nz ? amax * (1.f / 256.f) : 1.f;
    uint32_t t[4][3];
#pragma unroll
    for (int i = 0; i < 4; ++i) split3_e4m3x2(t[i], x[2 * i] * sq, x[2 * i + 1] * sq);
    const int chunk = d0 >> 4, within = d0 & 15;
#pragma unroll
    for (int term = 0; term < 4; ++term) {
      const int row = 32 * cg + 8 * term + h;
      uint2 v = make_uint2(0u, 0u);
      if (term < 3) {
        v.x = t[0][term] | (t[1][term] << 16);
        v.y = t[2][term] | (t[3][term] << 16);
      }
      *reinterpret_cast<uint2*>(qbuf + swz<128>(row, chunk) + within) = v;
    }
  }
}

// v[lane >> 3][lane & 7] for lane < 16, without a local-memory array index
__device__ __forceinline__ float tc5_pick16(const float (&v)[2][8], int lane) {
  float mine = v[0][0];
#pragma unroll
  for (int c = 0; c < 2; ++c)
#pragma unroll
    for (int j = 0; j < 8; ++j) mine = (lane & 15) == 8 * c + j ? v[c][j] : mine;
  return mine;
}

template <int QT, int KVF>
__device__ __forceinline__ void tc5_softmax_wide(const DecodeParams& p, const SegSource& src,
                                                 uint8_t* smem_ring, uint8_t* qbuf, uint8_t* pbuf,
                                                 float* scratch, float* pair_scratch, const Tc5Bars& bars,
                                                 uint32_t tmem, int num_stages) {
  using G = TileGeom<KVF, 128>;
  using TC = Tc5Cols<true>;
  constexpr int HD = 128;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int GS = p.group_size;  // <= 16 columns
  const int hq_real = p.num_qo_heads / p.q_rows;
  const bool use_cap = p.soft_cap > 0.f;
  const float cap_pre = use_cap ? p.sm_scale / p.soft_cap : 0.f;
  const float cap_post = use_cap ? p.soft_cap * 1.4426950408889634f : 0.f;
  const uint32_t lane_base = (uint32_t)(warp * 32) << 16;  // TMEM lane field
  // scratch (floats): wmax [4][16] @0, dq [16] @72, tmem slot @120, vote [2][4] @124;
  // pair_scratch: lsum [4][16] @0, per column: causal cut (int) @64, ALiBi slope @80
  float* wmax = scratch;
  float* dqs = scratch + 72;
  float* lsum = pair_scratch;
  int* hid_sm = reinterpret_cast<int*>(pair_scratch + 64);
  float* slope_sm = pair_scratch + 80;
  volatile int* vote = reinterpret_cast<volatile int*>(scratch + 124);

  int stage = 0;
  uint32_t n = 0;  // tiles consumed so far: S buffer = n & 1
  PendingMerge pm;
  pm.mu = -1; pm.old = pm.cnt = pm.slot = pm.req = pm.head = 0;
  bool have_q = false;
  for (int k = 0;; ++k) {
    Segment sg;
    if (!seg_fetch(src, k, sg)) break;
    if (!have_q) {
      if (k > 0) mbar_wait(bars.q_free, (k - 1) & 1u);
      tc5_prepare_q_wide<QT>(p, sg, qbuf, dqs, tid);
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(bars.q_ready[0]);
    }
    if (tid == 0) {
      src.ctl[kQPeek] = seg_peek(src, k + 1) ? 1 : 0;
      src.ctl[kQDone0] = k;
    }
    if (tid < 16) {
      const RowHead rh = row_head(p, sg.kv_head, tid);
      hid_sm[tid] = p.q_rows_noncausal ? 0 : p.q_rows - 1 - rh.row;
      float a = 0.f;
      if (p.alibi_slopes != nullptr && tid < GS)
        a = __ldg(p.alibi_slopes + rh.head) * (use_cap ? cap_pre : 1.4426950408889634f);
      slope_sm[tid] = a;
    }
    named_bar_sync(1, 128);
    have_q = src.ctl[kQPeek] != 0;
    const float pre_scale = p.alibi_slopes != nullptr ? cap_pre * p.sm_scale : cap_pre;
    float rho[2][8], l[2][8];
#pragma unroll
    for (int c = 0; c < 2; ++c)
#pragma unroll
      for (int j = 0; j < 8; ++j) { rho[c][j] = kNegInf; l[c][j] = 0.f; }

    for (int tile = sg.tile_begin; tile < sg.tile_end; ++tile) {
      const uint32_t b = n & 1u;
      const int tok = tile * G::TT + tid;
      const bool inwin = tok >= sg.tok_lo && tok < sg.tok_hi;
      float x[2][8];
      bool vcol[2][8];
      mbar_wait(bars.s_ready[b], (n >> 1) & 1u);
      tc5::fence_after_sync();
      {
        const uint32_t sb = tmem + lane_base + kTc5ColS + TC::kSStride * b;
        float c0[16], c1[8], e0[16], e1[8];
        tc5::ld_x16(sb, c0);
        tc5::ld_x8(sb + 16u, c1);
        tc5::ld_x16(sb + 32u, e0);
        tc5::ld_x8(sb + 48u, e1);
        tc5::wait_ld();
        tc5::fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(bars.s_free[b]);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          x[0][j] = c0[j] + c0[8 + j] + c1[j];
          x[1][j] = e0[j] + e0[8 + j] + e1[j];
        }
      }
      bool over = false;
#pragma unroll
      for (int c = 0; c < 2; ++c)
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int col = 8 * c + j;
          const float d = dqs[col];
          const float v = use_cap ? cap_post * tanhf_approx(x[c][j] * (d * pre_scale) + slope_sm[col] * (float)tok)
                                  : x[c][j] * (d * p.sm_scale_log2) + slope_sm[col] * (float)tok;
          vcol[c][j] = inwin && tok < sg.tok_hi - hid_sm[col];
          x[c][j] = vcol[c][j] ? v : kNegInf;
          over |= x[c][j] > rho[c][j] + kTc5Bump;
        }
      const bool wover = __any_sync(0xffffffffu, over);
      if (lane == 0) vote[b * 4 + warp] = wover ? 1 : 0;
      named_bar_sync(2, 128);
      const bool bump = (vote[b * 4] | vote[b * 4 + 1] | vote[b * 4 + 2] | vote[b * 4 + 3]) != 0;
      if (bump) {  // CTA-uniform
        float mt[2][8];
#pragma unroll
        for (int c = 0; c < 2; ++c)
#pragma unroll
          for (int j = 0; j < 8; ++j) mt[c][j] = tc5_head_max<false>(x[c][j]);
        if (lane < 16) wmax[warp * 16 + lane] = tc5_pick16(mt, lane);
        named_bar_sync(3, 128);
        float alpha[2][8];
#pragma unroll
        for (int c = 0; c < 2; ++c)
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const int i = 8 * c + j;
            const float m = fmaxf(fmaxf(wmax[i], wmax[16 + i]), fmaxf(wmax[32 + i], wmax[48 + i]));
            const float rn = m > rho[c][j] + kTc5Bump ? m : rho[c][j];
            alpha[c][j] = ex2f(rho[c][j] - rn);
            l[c][j] *= alpha[c][j];
            rho[c][j] = rn;
          }
        if (tile != sg.tile_begin) {  // O is relative to the old rho: wait for P.V of the previous tile, rescale
          mbar_wait(bars.p_free[0], (n - 1) & 1u);
          tc5::fence_after_sync();
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            float o[16];
            tc5::ld_x16(tmem + lane_base + TC::kO + 16u * c, o);
            tc5::wait_ld();
#pragma unroll
            for (int j = 0; j < 8; ++j) { o[j] *= alpha[c][j]; o[8 + j] *= alpha[c][j]; }
            tc5::st_x16(tmem + lane_base + TC::kO + 16u * c, o);
          }
          tc5::wait_st();
          tc5::fence_before_sync();
        }
      }
      if (n > 0) mbar_wait(bars.p_free[0], (n - 1) & 1u);  // P.V of the previous tile is done with the buffer
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float p0 = vcol[c][2 * i] ? ex2f(x[c][2 * i] - rho[c][2 * i] + kTc5PShift) : 0.f;
          const float p1 = vcol[c][2 * i + 1] ? ex2f(x[c][2 * i + 1] - rho[c][2 * i + 1] + kTc5PShift) : 0.f;
          l[c][2 * i] += p0;
          l[c][2 * i + 1] += p1;
          w[i] = split_e4m3x2(p0, p1);  // hi pair in bits 0-15, lo pair in bits 16-31
        }
        uint4 row;
        row.x = (w[0] & 0xffffu) | (w[1] << 16);
        row.y = (w[2] & 0xffffu) | (w[3] << 16);
        row.z = (w[0] >> 16) | (w[1] & 0xffff0000u);
        row.w = (w[2] >> 16) | (w[3] & 0xffff0000u);
        *reinterpret_cast<uint4*>(pbuf + c * 2048 + tid * 16) = row;
      }
      if (tok >= sg.tok_hi) {  // this token's V row was not (validly) loaded: 0 * junk must not be NaN
        uint4* vrow = reinterpret_cast<uint4*>(smem_ring + stage * G::STAGE_BYTES + G::KV_BYTES + tid * 128);
#pragma unroll
        for (int i = 0; i < 8; ++i) vrow[i] = make_uint4(0u, 0u, 0u, 0u);
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(bars.p_ready[0]);
      ++n;
      if (++stage == num_stages) stage = 0;
    }
    if (have_q) {  // CTA-uniform: the next segment's query, once every Q.K MMA of this one is complete
      if (sg.tile_end <= sg.tile_begin) named_bar_sync(2, 128);  // dq of this segment is still being read
      mbar_wait(bars.q_free, k & 1u);
      const Segment sn = seg_read(src, k + 1);
      tc5_prepare_q_wide<QT>(p, sn, qbuf, dqs, tid);
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(bars.q_ready[0]);
    }

    // ---- segment epilogue: thread tid <-> head-dim index d = tid (TMEM lane of O) ---------------
    float o8[16];
    if (sg.tile_end > sg.tile_begin) {
      mbar_wait(bars.p_free[0], (n - 1) & 1u);  // P.V of the last tile complete
      tc5::fence_after_sync();
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        float o[16];
        tc5::ld_x16(tmem + lane_base + TC::kO + 16u * c, o);
        tc5::wait_ld();
#pragma unroll
        for (int j = 0; j < 8; ++j) o8[8 * c + j] = o[j] + o[8 + j];
      }
      tc5::fence_before_sync();
    } else {
#pragma unroll
      for (int i = 0; i < 16; ++i) o8[i] = 0.f;
    }
    {
      float L[2][8];
#pragma unroll
      for (int c = 0; c < 2; ++c)
#pragma unroll
        for (int j = 0; j < 8; ++j) L[c][j] = tc5_head_sum<false>(l[c][j]);
      if (lane < 16) lsum[warp * 16 + lane] = tc5_pick16(L, lane);
    }
    if (tid == 0) merge_post(p, pm, scratch);
    named_bar_sync(3, 128);
    merge_run_posted<QT, HD, true>(p, scratch, tid);

    const bool direct = sg.slot < 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      if (i < GS) {
        const RowHead rh = row_head(p, sg.kv_head, i);
        float Lj = (lsum[i] + lsum[16 + i]) + (lsum[32 + i] + lsum[48 + i]);
        const float rj = rho[i >> 3][i & 7];
        const bool any = Lj > 0.f;
        if (direct && any && p.sinks != nullptr)
          Lj += ex2f(__ldg(p.sinks + rh.head) * 1.4426950408889634f - rj + kTc5PShift);
        const float val = any ? o8[i] * ((direct ? p.v_scale : 1.f) / Lj) : 0.f;
        const float lse2 = any ? rj + lg2f(Lj) - kTc5PShift : kNegInf;
        if (direct) {
          const int64_t oh = rh.token(p, sg.request) * hq_real + rh.head;
          if (QT == 1) static_cast<__nv_bfloat16*>(p.o)[oh * HD + tid] = __float2bfloat16_rn(val);
          else static_cast<__half*>(p.o)[oh * HD + tid] = __float2half_rn(val);
          if (p.lse != nullptr && tid == i) p.lse[oh] = lse2;
        } else {
          p.tmp_o[((int64_t)sg.slot * GS + i) * HD + tid] = val;
          if (tid == i) p.tmp_lse[(int64_t)sg.slot * GS + i] = lse2;
        }
      }
    }
    if (sg.merge_unit >= 0) {
      named_bar_sync(2, 128);
      if (tid == 0) merge_arrive(p, pm, sg);
    }
  }
  if (tid == 0) merge_post(p, pm, scratch);
  named_bar_sync(3, 128);
  merge_run_posted<QT, HD, true>(p, scratch, tid);
}

// QT: 0 f16, 1 bf16 (q / o dtype).  KVF: 1 = e4m3, 2 = e5m2 KV.  head_dim 128, TMA-eligible pages.
// HP: head-pair plan (see the top of this file).  HD: 128, or 256 (one CTA per SM: 64 KiB stages, two K / V /
// Qs panels per row, two O accumulators; single-head units only).
template <int QT, int KVF, int MODE = kTc5Single, int HD = 128>
__global__ void __launch_bounds__(kTc5Threads, HD == 128 ? 2 : 1)
    paged_decode_tc5_kernel(const __grid_constant__ DecodeParams p,
                            const __grid_constant__ CUtensorMap kmap,
                            const __grid_constant__ CUtensorMap vmap) {
  using G = TileGeom<KVF, HD>;
  using L = SmemLayout<KVF, HD>;
  constexpr bool HP = MODE == kTc5Pairs, WIDE = MODE == kTc5Wide, Q1 = MODE != kTc5Single;
  static_assert(L::SCRATCH_O >= 2 * 32 * HD + 2 * 2048, "Q / P buffers do not fit the scratch region");
  static_assert(MODE == kTc5Single || HD == 128, "head pairs / wide units: head_dim 128");
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t pad = ((raw + 1023u) & ~1023u) - raw;
  uint8_t* base = smem_raw + pad;
  const uint32_t ring = raw + pad;
  const int ring_bytes = p.num_stages * G::STAGE_BYTES;
  uint8_t* qbufs = base + ring_bytes;            // 2 x 4 KiB (HP: 1 x 8 KiB), 1024-aligned
  uint8_t* pbufs = qbufs + 2 * 32 * HD;          // 2 x 2 KiB
  float* scratch = reinterpret_cast<float*>(base + ring_bytes + L::SCRATCH_O);
  Segment* seg_cache =
      reinterpret_cast<Segment*>(base + ring_bytes + L::SCRATCH_O + L::SCRATCH_ML);
  const uint32_t full_bar = ring + ring_bytes + L::SCRATCH_O + L::SCRATCH_ML + L::SEG_BYTES;
  const uint32_t empty_bar = full_bar + 8 * kMaxStages;
  const Tc5Bars bars = tc5_bars(empty_bar + 8 * kMaxStages);
  // HP only (the launch adds kTc5PairScratchBytes): right after the layout every kernel shares
  float* pair_scratch = reinterpret_cast<float*>(base + ring_bytes + L::SCRATCH_O + L::SCRATCH_ML +
                                                 L::SEG_BYTES + L::BAR_BYTES);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(scratch + 120);

  const int tid = threadIdx.x, warp = tid >> 5;
  FI_TRACE_IF(tid == 0, 0);
  // TMEM allocation (warp 1) overlaps barrier setup (thread 0) and the work-list fetch
  if (warp == 1) tc5::tmem_alloc(smem_u32(tmem_slot), Tc5Cols<Q1>::kAlloc);
  if (tid == 0) {
    for (int s = 0; s < p.num_stages; ++s) {
      mbar_init(full_bar + 8 * s, 1);
      mbar_init(empty_bar + 8 * s, 1);  // one tcgen05.commit per tile
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(bars.s_ready[i], 1);
      mbar_init(bars.s_free[i], 4);
      mbar_init(bars.p_ready[i], 4);
      mbar_init(bars.p_free[i], 1);
      mbar_init(bars.q_ready[i], 4);
    }
    mbar_init(bars.q_free, 1);
    fence_mbar_init();
    tma_prefetch_desc(&kmap);
    tma_prefetch_desc(&vmap);
  }
  const SegSource src = seg_source_init(p, HP ? G::TT / 2 : G::TT, seg_cache, scratch, /*has_reader1=*/true);  // TT: 128 tokens at either head_dim
  tc5::fence_before_sync();
  __syncthreads();
  tc5::fence_after_sync();
  const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(tmem_slot);
  FI_TRACE_IF(tid == 0, 1);
  if (seg_source_idle(src)) {
    if (warp == 1) tc5::tmem_dealloc(tmem, Tc5Cols<Q1>::kAlloc);
    return;
  }
  if (p.enable_pdl == 1) pdl_launch_dependents();  // early trigger, see paged_decode_kernel
  FI_TRACE_IF(tid == 0, 2);

  // dependency wait per role (see paged_decode_kernel): the softmax warps read q, the producer
  // requests its first page indices and then waits, the MMA issuer touches no global memory
  if (warp < 4) {
    if (p.enable_pdl) pdl_wait();
    if constexpr (WIDE) tc5_softmax_wide<QT, KVF>(p, src, base, qbufs, pbufs, scratch, pair_scratch, bars, tmem, p.num_stages);
    else if constexpr (HP) tc5_softmax<QT, KVF, HP>(p, src, base, qbufs, pbufs, scratch, pair_scratch, bars, tmem, p.num_stages);
    else tc5_softmax_single<QT, KVF, HD>(p, src, base, qbufs, pbufs, scratch, bars, tmem, p.num_stages);
  } else if (warp == 4) {
    produce_tma<KVF, HD, false, HP>(p, src, &kmap, &vmap, ring, full_bar, empty_bar);
  } else {
    tc5_issue<KVF, MODE, HD>(p, src, ring, smem_u32(qbufs), smem_u32(pbufs), full_bar, empty_bar, bars, tmem);
  }
  if (p.enable_pdl == 2) pdl_launch_dependents();
  tc5::fence_before_sync();
  __syncthreads();
  if (warp == 1) tc5::tmem_dealloc(tmem, Tc5Cols<Q1>::kAlloc);
}

}  // namespace fib200
