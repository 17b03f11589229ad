// decode_mma.cu — instantiates and launches the mma.sync decode kernels (decode_kernel.cuh).
// Built as four objects, one per (q dtype, KV width) slice: build.py passes -DFI_SLICE_QT=0|1 and
// -DFI_SLICE_FP8=0|1; without them the file instantiates all four.  The loader / head_dim / variant
// choice follows BatchDecodeWithPagedKVCacheDispatched (reference
// include/flashinfer/attention/decode.cuh:739-825), which picks its kernel from (HEAD_DIM, GROUP_SIZE,
// POS_ENCODING_MODE) the same way.
#include "decode_kernel.cuh"
#include "decode_launch.h"

namespace fib200 {

template <int QT, int KVF, int HD, int LOADER, bool ROPE = false, int MROWS = 0>
static int launch_one(const DecodeParams& p, const CUtensorMap& km, const CUtensorMap& vm,
                      int grid, cudaStream_t stream) {
  auto kern = paged_decode_kernel<QT, KVF, HD, LOADER, ROPE, MROWS>;
  const int smem = SmemLayout<KVF, HD>::total(p.num_stages);
  FI_CUDA_TRY(ensure_smem_attr(kern, smem));
  return launch_pdl(kern, grid, BlockDim<LOADER>::kThreads, smem, p.enable_pdl != 0, stream, p, km, vm);
}

template <int QT, int KVF, int HD>
static int launch_loader(const DecodeParams& p, const CUtensorMap& km, const CUtensorMap& vm,
                         int grid, bool tma, cudaStream_t stream) {
  if constexpr (KVF == 0) {
    if (p.q_rows >= 1 && p.group_size > kMaxGroup) {  // wide units: two N tiles per warp (TMA loader, D <= 128)
      if constexpr (HD <= 128) {
        if (tma) return launch_one<QT, KVF, HD, 0, false, 2>(p, km, vm, grid, stream);
      }
      return fail(FI_ERR_REPLAN, "the plan was made with FI_PLAN_WIDE_UNITS, but this run() needs the gather loader");
    }
    if (p.q_rows > 1)  // multi-row units (fi_batch_decode_run_args q_rows): checked by the caller
      return tma ? launch_one<QT, KVF, HD, 0, false, 1>(p, km, vm, grid, stream)
                 : launch_one<QT, KVF, HD, 1, false, 1>(p, km, vm, grid, stream);
    // ROPE_LLAMA on 16-bit KV: K rotated in registers (decode_kernel.cuh)
    if (p.rope)
      return tma ? launch_one<QT, KVF, HD, 0, true>(p, km, vm, grid, stream)
                 : launch_one<QT, KVF, HD, 1, true>(p, km, vm, grid, stream);
  }
  return tma ? launch_one<QT, KVF, HD, 0>(p, km, vm, grid, stream)
             : launch_one<QT, KVF, HD, 1>(p, km, vm, grid, stream);
}

template <int QT, int KVF>
static int launch_hd(const DecodeParams& p, const CUtensorMap& km, const CUtensorMap& vm,
                     int head_dim, int grid, bool tma, cudaStream_t stream) {
  switch (head_dim) {
    case 128: return launch_loader<QT, KVF, 128>(p, km, vm, grid, tma, stream);
    case 256: return launch_loader<QT, KVF, 256>(p, km, vm, grid, tma, stream);
    case 64:
      if constexpr (KVF == 0) return launch_loader<QT, KVF, 64>(p, km, vm, grid, tma, stream);
    default: break;
  }
  return fail(FI_ERR_UNSUPPORTED, "unsupported head_dim %d", head_dim);
}

template <int QT, bool FP8>
int launch_mma_slice(const DecodeParams& p, const CUtensorMap& km, const CUtensorMap& vm, int kvf,
                     int head_dim, int grid, bool tma, cudaStream_t stream) {
  if constexpr (!FP8) {
    if (kvf == 0) return launch_hd<QT, 0>(p, km, vm, head_dim, grid, tma, stream);
  } else {
    if (kvf == 1) return launch_hd<QT, 1>(p, km, vm, head_dim, grid, tma, stream);
    if (kvf == 2) return launch_hd<QT, 2>(p, km, vm, head_dim, grid, tma, stream);
  }
  return fail(FI_ERR_UNSUPPORTED, "unsupported kv format");
}

#if !defined(FI_SLICE_QT) || (FI_SLICE_QT == 0 && FI_SLICE_FP8 == 0)
template int launch_mma_slice<0, false>(const DecodeParams&, const CUtensorMap&, const CUtensorMap&, int, int, int, bool, cudaStream_t);
#endif
#if !defined(FI_SLICE_QT) || (FI_SLICE_QT == 1 && FI_SLICE_FP8 == 0)
template int launch_mma_slice<1, false>(const DecodeParams&, const CUtensorMap&, const CUtensorMap&, int, int, int, bool, cudaStream_t);
#endif
#if !defined(FI_SLICE_QT) || (FI_SLICE_QT == 0 && FI_SLICE_FP8 == 1)
template int launch_mma_slice<0, true>(const DecodeParams&, const CUtensorMap&, const CUtensorMap&, int, int, int, bool, cudaStream_t);
#endif
#if !defined(FI_SLICE_QT) || (FI_SLICE_QT == 1 && FI_SLICE_FP8 == 1)
template int launch_mma_slice<1, true>(const DecodeParams&, const CUtensorMap&, const CUtensorMap&, int, int, int, bool, cudaStream_t);
#endif

}  // namespace fib200
