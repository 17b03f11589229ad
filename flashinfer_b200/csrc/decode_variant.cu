// decode_variant.cu — instantiates and launches the CUDA-core variant kernel (decode_rope_kernel.cuh):
// pos_encoding_mode = ROPE_LLAMA with fp8 KV, fp8 KV at head_dim 64, head_dim 512.  Built as two objects,
// one per q dtype: build.py passes -DFI_SLICE_QT=0|1; without it the file instantiates both.
#include "decode_launch.h"
#include "decode_rope_kernel.cuh"

namespace fib200 {

template <int QT, int KVF>
static int launch_rope(const DecodeParams& p, int head_dim, int grid, cudaStream_t stream) {
  const int block = rope_warps(head_dim) * 32;
  const bool pdl = p.enable_pdl != 0;
#define FI_ROPE_CASE(HD_)                                                                         \
  case HD_:                                                                                       \
    return p.group_size <= 4                                                                      \
               ? launch_pdl(paged_decode_rope_kernel<QT, KVF, HD_, 4>, grid, block, 0, pdl, stream, p) \
               : launch_pdl(paged_decode_rope_kernel<QT, KVF, HD_, 8>, grid, block, 0, pdl, stream, p);
  switch (head_dim) {
    FI_ROPE_CASE(64)
    FI_ROPE_CASE(128)
    FI_ROPE_CASE(256)
    case 512:  // groups above 4 were split into virtual KV heads by the planner (max_group_for)
      if (p.group_size > 4) return fail(FI_ERR_UNSUPPORTED, "head_dim 512 takes GQA groups of at most 4 per unit");
      return launch_pdl(paged_decode_rope_kernel<QT, KVF, 512, 4>, grid, block, 0, pdl, stream, p);
    default: break;
  }
#undef FI_ROPE_CASE
  return fail(FI_ERR_UNSUPPORTED, "unsupported head_dim %d", head_dim);
}

template <int QT>
int launch_variant_slice(const DecodeParams& p, int kvf, int head_dim, int grid, cudaStream_t stream) {
  switch (kvf) {
    case 0: return launch_rope<QT, 0>(p, head_dim, grid, stream);
    case 1: return launch_rope<QT, 1>(p, head_dim, grid, stream);
    case 2: return launch_rope<QT, 2>(p, head_dim, grid, stream);
    default: break;
  }
  return fail(FI_ERR_UNSUPPORTED, "unsupported kv format");
}

#if !defined(FI_SLICE_QT) || FI_SLICE_QT == 0
template int launch_variant_slice<0>(const DecodeParams&, int, int, int, cudaStream_t);
#endif
#if !defined(FI_SLICE_QT) || FI_SLICE_QT == 1
template int launch_variant_slice<1>(const DecodeParams&, int, int, int, cudaStream_t);
#endif

}  // namespace fib200
