// decode_launch.h — the seam between decode.cu (host logic, C-ABI) and the translation units that
// instantiate the kernels.  The 114 decode kernels are independent templates, so they compile in
// parallel as slices (build.py: one object per (family, q dtype[, KV width]) slice) instead of one
// 6-minute translation unit; decode.cu only sees these declarations.
//   decode_mma.cu      launch_mma_slice<QT, FP8>   mma.sync kernels        (decode_kernel.cuh)
//   decode_tc5.cu      launch_tc5_slice<QT>        tcgen05 / TMEM kernels  (decode_kernel_tc5.cuh)
//   decode_variant.cu  launch_variant_slice<QT>    CUDA-core variant kernel (decode_rope_kernel.cuh)
// QT: 0 = fp16 q/o, 1 = bf16 q/o.  The reference reaches the same split through its JIT: one module
// per (dtype, head_dim, variant) (flashinfer/jit/attention/modules.py gen_batch_decode_module).
#pragma once

#include <cuda.h>
#include <cuda_runtime.h>

#include <map>
#include <utility>

#include "decode_params.h"
#include "host_common.h"

namespace fib200 {

template <int QT, bool FP8>
int launch_mma_slice(const DecodeParams& p, const CUtensorMap& km, const CUtensorMap& vm, int kvf,
                     int head_dim, int grid, bool tma, cudaStream_t stream);
template <int QT>
int launch_tc5_slice(const DecodeParams& p, const CUtensorMap& km, const CUtensorMap& vm, int kvf,
                     int mode, int head_dim, int grid, cudaStream_t stream);
template <int QT>
int launch_variant_slice(const DecodeParams& p, int kvf, int head_dim, int grid, cudaStream_t stream);

// The opt-in dynamic shared memory limit is a per-(function, device) attribute: set it when it
// has to grow (or the device changed), not on every launch.
template <typename K>
static cudaError_t ensure_smem_attr(K kern, int smem) {
  // all kernels share one pointer type, so the cache is keyed by (function, device)
  static thread_local std::map<std::pair<const void*, int>, int> cache;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  int& cur = cache[{reinterpret_cast<const void*>(kern), dev}];
  if (smem <= cur) return cudaSuccess;
  e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e == cudaSuccess) cur = smem;
  return e;
}

// one launch, optionally as a programmatic dependent launch (DecodeParams.enable_pdl)
template <typename K, typename... Args>
static int launch_pdl(K kern, int grid, int block, int smem, bool pdl, cudaStream_t stream,
                      const Args&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(block);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  FI_CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, args...));
  return FI_OK;
}

}  // namespace fib200
