// gather.cu — peer-write gather of the per-rank output heads (KV-head sharding over NVLink).
//
// No counterpart in the reference's decode path (frameworks on top of it all-gather O with NCCL;
// the reference's own peer-memory machinery lives in flashinfer/comm/cuda_ipc.py and
// flashinfer/comm/vllm_ar.py:105-155 and was read for the protocol, not the code).  After the
// decode kernel of rank r has written o_local [rows, row_bytes] (its Hq / N heads of every
// request), ONE kernel — chained to the decode launch with programmatic dependent launch — stores
// every 16-byte chunk into the [rows, N * row_bytes] buffer of EVERY rank at rank r's head offset:
//   * through the NVLS multicast address of the symmetric buffer when there is one
//     (multimem.st: one store instruction, the NVSwitch replicates it to all N GPUs), else
//   * with one peer-mapped st.global per rank (P2P over NVLink);
// the CTA that finishes last then raises flag[r] = epoch on every rank (st.release.sys after a
// system fence) and waits until all N flags of its own rank carry the epoch (ld.acquire.sys), so when
// the kernel completes on a rank, that rank's full buffer holds every rank's heads.
// Epochs only grow; the destination is double-buffered by epoch parity on the Python side
// (flashinfer_b200/sharding.py PeerGather), which is what makes re-use safe without a second barrier.
#include <string.h>

#include "host_common.h"
#include "ptx.cuh"

namespace fib200 {

constexpr int kMaxPeers = 16;
// layout of a rank's control block (uint32 words, symmetric memory, zero at creation)
constexpr int kCtlFlags = 0;     // [kMaxPeers] flag[p] = last epoch rank p finished writing to us
constexpr int kCtlCounter = 32;  // CTA completion counter of the local launch (returns to zero)
constexpr int kCtlStatus = 33;   // != 0: a wait timed out (a peer never arrived)

struct PeerGatherParams {
  const char* src;
  char* dst[kMaxPeers];       // unicast: rank p's destination buffer as mapped here
  char* mc_dst;               // multicast address of the destination buffer, or null
  uint32_t* ctl[kMaxPeers];   // rank p's control block as mapped here
  int64_t rows, row_bytes, dst_row_stride, dst_offset;
  int32_t world, rank;
  uint32_t epoch;
  int32_t enable_pdl;
  unsigned long long timeout_ns;
};

__device__ __forceinline__ void st_release_sys_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st_v4(char* mc, const int4& v) {
  // .v4.f32 moves the 16 bytes as they are (a store does no arithmetic)
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc),
               "f"(__int_as_float(v.x)), "f"(__int_as_float(v.y)), "f"(__int_as_float(v.z)),
               "f"(__int_as_float(v.w))
               : "memory");
}
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// UNIT: the widest type every address and the row length allow (int4 / uint32_t / uint16_t)
template <typename UNIT>
__global__ void __launch_bounds__(1024) peer_gather_kernel(const PeerGatherParams p) {
  __shared__ int s_last;
  if (p.enable_pdl) pdl_wait();  // o_local is the previous kernel's output
  const int64_t upr = p.row_bytes / (int64_t)sizeof(UNIT);
  const int64_t total = p.rows * upr;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t row = i / upr, c = i - row * upr;
    const UNIT v = *reinterpret_cast<const UNIT*>(p.src + row * p.row_bytes + c * sizeof(UNIT));
    const int64_t off = p.dst_offset + row * p.dst_row_stride + c * (int64_t)sizeof(UNIT);
    if constexpr (sizeof(UNIT) == 16) {
      if (p.mc_dst != nullptr) {
        multimem_st_v4(p.mc_dst + off, v);
        continue;
      }
    }
    for (int d = 0; d < p.world; ++d) {  // start at our own rank: the ranks hit different peers first
      const int peer = (p.rank + d) % p.world;
      *reinterpret_cast<UNIT*>(p.dst[peer] + off) = v;
    }
  }
  // this CTA's stores are performed system-wide before its arrival is counted
  __threadfence_system();
  __syncthreads();
  uint32_t* my = p.ctl[p.rank];
  if (gridDim.x > 1) {  // small payloads run as ONE CTA: no counter round trip, no second fence
    if (threadIdx.x == 0)
      s_last = atom_add_acq_rel_gpu(reinterpret_cast<int*>(my + kCtlCounter), 1) == (int)gridDim.x - 1;
    __syncthreads();
    if (!s_last) return;
    if (threadIdx.x == 0) my[kCtlCounter] = 0;  // re-armed for the next launch
    __threadfence_system();
  }
  if ((int)threadIdx.x < p.world) {
    const int peer = threadIdx.x;
    st_release_sys_u32(p.ctl[peer] + kCtlFlags + p.rank, p.epoch);
    const unsigned long long t0 = global_ns();
    // (int32) difference: correct across the 2^32 wrap of the epoch counter
    while ((int32_t)(ld_acquire_sys_u32(my + kCtlFlags + peer) - p.epoch) < 0) {
      __nanosleep(100);
      if (global_ns() - t0 > p.timeout_ns) {
        my[kCtlStatus] = 1u + (uint32_t)peer;
        break;
      }
    }
  }
}

}  // namespace fib200

using namespace fib200;

extern "C" int fi_peer_gather(const void* src, void* const* peer_dst, void* mc_dst,
                              void* const* peer_ctl, int world, int rank, int64_t rows,
                              int64_t row_bytes, int64_t dst_row_stride_bytes,
                              int64_t dst_offset_bytes, uint32_t epoch, int enable_pdl, int device,
                              void* stream) {
  if (world < 1 || world > kMaxPeers || rank < 0 || rank >= world)
    return fail(FI_ERR_INVALID_ARGUMENT, "fi_peer_gather: bad rank %d / world %d (max %d)", rank, world,
                kMaxPeers);
  if (rows < 0 || row_bytes <= 0 || dst_row_stride_bytes < row_bytes || dst_offset_bytes < 0)
    return fail(FI_ERR_INVALID_ARGUMENT, "fi_peer_gather: bad geometry");
  if (src == nullptr || peer_dst == nullptr || peer_ctl == nullptr)
    return fail(FI_ERR_INVALID_ARGUMENT, "fi_peer_gather: null pointer");
  PeerGatherParams p;
  memset(&p, 0, sizeof(p));
  p.src = static_cast<const char*>(src);
  uintptr_t align = reinterpret_cast<uintptr_t>(src) | (uintptr_t)row_bytes |
                    (uintptr_t)dst_row_stride_bytes | (uintptr_t)dst_offset_bytes;
  for (int i = 0; i < world; ++i) {
    if (peer_dst[i] == nullptr || peer_ctl[i] == nullptr)
      return fail(FI_ERR_INVALID_ARGUMENT, "fi_peer_gather: null peer pointer %d", i);
    p.dst[i] = static_cast<char*>(peer_dst[i]);
    p.ctl[i] = static_cast<uint32_t*>(peer_ctl[i]);
    align |= reinterpret_cast<uintptr_t>(peer_dst[i]);
  }
  p.mc_dst = static_cast<char*>(mc_dst);
  align |= reinterpret_cast<uintptr_t>(mc_dst);
  p.rows = rows; p.row_bytes = row_bytes;
  p.dst_row_stride = dst_row_stride_bytes; p.dst_offset = dst_offset_bytes;
  p.world = world; p.rank = rank; p.epoch = epoch; p.enable_pdl = enable_pdl;
  p.timeout_ns = 2000000000ull;  // 2 s: a peer that never launches must not hang the GPU
  if (align & 1) return fail(FI_ERR_INVALID_ARGUMENT, "fi_peer_gather: odd addresses / sizes");
  FI_CUDA_TRY(cudaSetDevice(device));
  if (rows == 0) return FI_OK;
  const int unit = (align & 15) == 0 ? 16 : ((align & 3) == 0 ? 4 : 2);
  if (unit != 16) p.mc_dst = nullptr;
  const int64_t total = rows * (row_bytes / unit);
  // up to 8192 units (128 KiB of 16-byte chunks: the decode outputs of C2 / C3 / C5 at 4 and 8 ranks)
  // one CTA of 1024 threads does it all and signals without the CTA counter; larger payloads spread
  // over up to 64 CTAs of 256 threads
  const bool one_cta = total <= 8192;
  int grid = one_cta ? 1 : (int)((total + 255) / 256);
  if (grid > 64) grid = 64;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(one_cta ? 1024 : 256);
  cfg.stream = static_cast<cudaStream_t>(stream);
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = enable_pdl ? 1 : 0;
  if (unit == 16) FI_CUDA_TRY(cudaLaunchKernelEx(&cfg, peer_gather_kernel<int4>, p));
  else if (unit == 4) FI_CUDA_TRY(cudaLaunchKernelEx(&cfg, peer_gather_kernel<uint32_t>, p));
  else FI_CUDA_TRY(cudaLaunchKernelEx(&cfg, peer_gather_kernel<uint16_t>, p));
  return FI_OK;
}
