// page.cu — the step right before decode in every serving iteration: scatter the new tokens'
// K / V rows into the paged cache, and the ragged-indptr -> (batch index, position) expansion that
// feeds it.  Replaces append_paged_kv_cache (reference csrc/page.cu:24-110 ->
// AppendPagedKVCacheKernel / AppendPagedKVCache, include/flashinfer/page.cuh:254-287,370-420) and the
// Triton get_batch_indices_positions_kernel (flashinfer/triton/page.py:21-42).
//
// Both are pure byte movers (HBM-bound: bytes = 2 x nnz x Hkv x D x sizeof(kv) read + written, plus
// the index arrays).  The append kernel is dtype-agnostic: a (token, head) row is D x elem_size
// bytes and is copied in the widest vector (16 / 8 / 4 / 2 / 1 bytes) that every base pointer and
// stride allows; consecutive threads take consecutive chunks of a row, then the next head, then
// the next token, so both the ragged source and the destination page rows are touched in full
// 128-byte lines.  The grid is sized in multiples of the SM count and strides over the chunks.
#include "host_common.h"

namespace fib200 {

struct AppendParams {
  const char* append_k;
  const char* append_v;
  int64_t ak_stride_n, ak_stride_h, av_stride_n, av_stride_h;  // bytes
  char* k_cache;
  char* v_cache;
  int64_t kv_stride_page, kv_stride_n, kv_stride_h;  // bytes
  const int32_t* batch_indices;
  const int32_t* positions;
  const int32_t* kv_indices;
  const int32_t* kv_indptr;
  int32_t page_size;
  int32_t num_heads;
  int32_t chunks_per_row;  // row bytes / VEC
  int64_t total_chunks;    // nnz * num_heads * chunks_per_row
};

template <int VEC>
struct VecType;
template <> struct VecType<16> { using T = uint4; };
template <> struct VecType<8> { using T = uint2; };
template <> struct VecType<4> { using T = uint32_t; };
template <> struct VecType<2> { using T = uint16_t; };
template <> struct VecType<1> { using T = uint8_t; };

template <int VEC>
__global__ void __launch_bounds__(256) append_paged_kv_kernel(const AppendParams p) {
  using T = typename VecType<VEC>::T;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t per_token = (int64_t)p.num_heads * p.chunks_per_row;
  for (int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; c < p.total_chunks; c += stride) {
    const int64_t i = c / per_token;
    const int rem = (int)(c - i * per_token);
    const int h = rem / p.chunks_per_row;
    const int ch = rem - h * p.chunks_per_row;
    // reference page.cuh:268-270: page_iter, entry = divmod(indptr[b] * page_size + pos, page_size)
    const int b = __ldg(p.batch_indices + i);
    const int pos = __ldg(p.positions + i);
    const int page_iter = __ldg(p.kv_indptr + b) + pos / p.page_size;
    const int entry = pos - (pos / p.page_size) * p.page_size;
    const int64_t page = __ldg(p.kv_indices + page_iter);
    const int64_t dst = page * p.kv_stride_page + (int64_t)entry * p.kv_stride_n +
                        (int64_t)h * p.kv_stride_h + (int64_t)ch * VEC;
    const T kv = __ldg(reinterpret_cast<const T*>(p.append_k + i * p.ak_stride_n + h * p.ak_stride_h +
                                                  (int64_t)ch * VEC));
    const T vv = __ldg(reinterpret_cast<const T*>(p.append_v + i * p.av_stride_n + h * p.av_stride_h +
                                                  (int64_t)ch * VEC));
    *reinterpret_cast<T*>(p.k_cache + dst) = kv;
    *reinterpret_cast<T*>(p.v_cache + dst) = vv;
  }
}

// one CTA per request (grid-stride over requests), threads stride over the request's new tokens:
// batch_indices[j] = i, positions[j] = j + seq_len[i] - append_indptr[i + 1]
__global__ void __launch_bounds__(128) batch_indices_positions_kernel(
    const int32_t* __restrict__ append_indptr, const int32_t* __restrict__ seq_lens,
    int32_t* __restrict__ batch_indices, int32_t* __restrict__ positions, int batch_size, int nnz) {
  for (int i = blockIdx.x; i < batch_size; i += gridDim.x) {
    const int begin = __ldg(append_indptr + i), end = __ldg(append_indptr + i + 1);
    const int off = __ldg(seq_lens + i) - end;
    for (int j = begin + threadIdx.x; j < end && j < nnz; j += blockDim.x) {
      batch_indices[j] = i;
      positions[j] = j + off;
    }
  }
}

static int widest_vec(uintptr_t bits, int row_bytes) {
  for (int v = 16; v > 1; v >>= 1)
    if ((bits & (uintptr_t)(v - 1)) == 0 && row_bytes % v == 0) return v;
  return 1;
}

}  // namespace fib200

using namespace fib200;

extern "C" {

int fi_get_batch_indices_positions(const int32_t* append_indptr, const int32_t* seq_lens,
                                   int32_t* batch_indices, int32_t* positions, int batch_size,
                                   int nnz, int device, void* stream) {
  if (batch_size < 0 || nnz < 0) return fail(FI_ERR_INVALID_ARGUMENT, "invalid batch_size / nnz");
  if (batch_size == 0 || nnz == 0) return FI_OK;
  if (append_indptr == nullptr || seq_lens == nullptr || batch_indices == nullptr || positions == nullptr)
    return fail(FI_ERR_INVALID_ARGUMENT, "null pointer");
  FI_CUDA_TRY(cudaSetDevice(device));
  int sms = 148;
  device_sm_count(device, &sms);
  const int grid = batch_size < sms * 8 ? batch_size : sms * 8;
  batch_indices_positions_kernel<<<grid, 128, 0, static_cast<cudaStream_t>(stream)>>>(
      append_indptr, seq_lens, batch_indices, positions, batch_size, nnz);
  FI_CUDA_TRY(cudaGetLastError());
  return FI_OK;
}

int fi_append_paged_kv_cache(const void* append_key, const void* append_value, int64_t ak_stride_n,
                             int64_t ak_stride_h, int64_t av_stride_n, int64_t av_stride_h,
                             const int32_t* batch_indices, const int32_t* positions, int64_t nnz,
                             void* k_cache, void* v_cache, int64_t kv_stride_page,
                             int64_t kv_stride_n, int64_t kv_stride_h, const int32_t* kv_indices,
                             const int32_t* kv_indptr, int num_kv_heads, int head_dim, int page_size,
                             int dtype, int device, void* stream) {
  const int es = dtype_size(dtype);
  if (es == 0) return fail(FI_ERR_UNSUPPORTED, "unsupported dtype %d", dtype);
  if (nnz < 0 || num_kv_heads <= 0 || head_dim <= 0 || page_size <= 0)
    return fail(FI_ERR_INVALID_ARGUMENT, "invalid append shape");
  if (nnz == 0) return FI_OK;
  if (!append_key || !append_value || !batch_indices || !positions || !k_cache || !v_cache ||
      !kv_indices || !kv_indptr)
    return fail(FI_ERR_INVALID_ARGUMENT, "null pointer");
  AppendParams p;
  p.append_k = static_cast<const char*>(append_key);
  p.append_v = static_cast<const char*>(append_value);
  p.ak_stride_n = ak_stride_n * es; p.ak_stride_h = ak_stride_h * es;
  p.av_stride_n = av_stride_n * es; p.av_stride_h = av_stride_h * es;
  p.k_cache = static_cast<char*>(k_cache);
  p.v_cache = static_cast<char*>(v_cache);
  p.kv_stride_page = kv_stride_page * es; p.kv_stride_n = kv_stride_n * es; p.kv_stride_h = kv_stride_h * es;
  p.batch_indices = batch_indices; p.positions = positions;
  p.kv_indices = kv_indices; p.kv_indptr = kv_indptr;
  p.page_size = page_size; p.num_heads = num_kv_heads;
  const int row_bytes = head_dim * es;
  const uintptr_t bits = reinterpret_cast<uintptr_t>(append_key) | reinterpret_cast<uintptr_t>(append_value) |
                         reinterpret_cast<uintptr_t>(k_cache) | reinterpret_cast<uintptr_t>(v_cache) |
                         (uintptr_t)p.ak_stride_n | (uintptr_t)p.ak_stride_h | (uintptr_t)p.av_stride_n |
                         (uintptr_t)p.av_stride_h | (uintptr_t)p.kv_stride_page |
                         (uintptr_t)p.kv_stride_n | (uintptr_t)p.kv_stride_h;
  const int vec = widest_vec(bits, row_bytes);
  p.chunks_per_row = row_bytes / vec;
  p.total_chunks = nnz * num_kv_heads * p.chunks_per_row;
  FI_CUDA_TRY(cudaSetDevice(device));
  int sms = 148;
  device_sm_count(device, &sms);
  int64_t blocks = (p.total_chunks + 255) / 256;
  const int64_t cap = (int64_t)sms * 8;  // grid-stride beyond 8 CTAs/SM
  if (blocks > cap) blocks = cap;
  cudaStream_t cs = static_cast<cudaStream_t>(stream);
  switch (vec) {
    case 16: append_paged_kv_kernel<16><<<(int)blocks, 256, 0, cs>>>(p); break;
    case 8: append_paged_kv_kernel<8><<<(int)blocks, 256, 0, cs>>>(p); break;
    case 4: append_paged_kv_kernel<4><<<(int)blocks, 256, 0, cs>>>(p); break;
    case 2: append_paged_kv_kernel<2><<<(int)blocks, 256, 0, cs>>>(p); break;
    default: append_paged_kv_kernel<1><<<(int)blocks, 256, 0, cs>>>(p); break;
  }
  FI_CUDA_TRY(cudaGetLastError());
  return FI_OK;
}

}  // extern "C"
