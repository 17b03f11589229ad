// decode_tc5.cu — instantiates and launches the tcgen05 / TMEM decode kernels for fp8 KV
// (decode_kernel_tc5.cuh).  Built as two objects, one per q dtype: build.py passes -DFI_SLICE_QT=0|1;
// without it the file instantiates both.
#include "decode_kernel_tc5.cuh"
#include "decode_launch.h"

namespace fib200 {

template <int QT, int KVF, int MODE = kTc5Single, int HD = 128>
static int launch_tc5(const DecodeParams& p, const CUtensorMap& km, const CUtensorMap& vm, int grid,
                      cudaStream_t stream) {
  auto kern = paged_decode_tc5_kernel<QT, KVF, MODE, HD>;
  const int smem = SmemLayout<KVF, HD>::total(p.num_stages) + (MODE != kTc5Single ? kTc5PairScratchBytes : 0);
  FI_CUDA_TRY(ensure_smem_attr(kern, smem));
  return launch_pdl(kern, grid, kTc5Threads, smem, p.enable_pdl != 0, stream, p, km, vm);
}

template <int QT, int KVF>
static int launch_tc5_mode(const DecodeParams& p, const CUtensorMap& km, const CUtensorMap& vm, int mode,
                           int head_dim, int grid, cudaStream_t stream) {
  if (head_dim == 128) {
    switch (mode) {
      case kTc5Single: return launch_tc5<QT, KVF>(p, km, vm, grid, stream);
      case kTc5Pairs: return launch_tc5<QT, KVF, kTc5Pairs>(p, km, vm, grid, stream);
      case kTc5Wide: return launch_tc5<QT, KVF, kTc5Wide>(p, km, vm, grid, stream);
      default: break;
    }
  } else if (head_dim == 256 && mode == kTc5Single) {
    return launch_tc5<QT, KVF, kTc5Single, 256>(p, km, vm, grid, stream);
  }
  return fail(FI_ERR_UNSUPPORTED, "no tcgen05 kernel for head_dim %d in mode %d", head_dim, mode);
}

template <int QT>
int launch_tc5_slice(const DecodeParams& p, const CUtensorMap& km, const CUtensorMap& vm, int kvf,
                     int mode, int head_dim, int grid, cudaStream_t stream) {
  if (kvf == 1) return launch_tc5_mode<QT, 1>(p, km, vm, mode, head_dim, grid, stream);
  if (kvf == 2) return launch_tc5_mode<QT, 2>(p, km, vm, mode, head_dim, grid, stream);
  return fail(FI_ERR_UNSUPPORTED, "the tcgen05 kernels take fp8 KV");
}

#if !defined(FI_SLICE_QT) || FI_SLICE_QT == 0
template int launch_tc5_slice<0>(const DecodeParams&, const CUtensorMap&, const CUtensorMap&, int, int, int, int, cudaStream_t);
#endif
#if !defined(FI_SLICE_QT) || FI_SLICE_QT == 1
template int launch_tc5_slice<1>(const DecodeParams&, const CUtensorMap&, const CUtensorMap&, int, int, int, int, cudaStream_t);
#endif

}  // namespace fib200
