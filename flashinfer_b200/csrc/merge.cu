// merge.cu — C-ABI entry points of the attention-state merge operators.
// Replaces merge_state / merge_state_in_place / merge_states (reference csrc/cascade.cu:23,59,102)
// and their launchers MergeState / MergeStateInPlace / MergeStates /
// VariableLengthMergeStates (include/flashinfer/attention/cascade.cuh:573,605,638,686).
#include "host_common.h"
#include "merge_kernels.cuh"

namespace fib200 {

static int check_merge_shape(int64_t seq_len, int num_heads, int head_dim, int dtype) {
  if (seq_len < 0 || num_heads <= 0 || head_dim <= 0)
    return fail(FI_ERR_INVALID_ARGUMENT, "invalid merge shape");
  if (head_dim % 8 != 0)
    return fail(FI_ERR_UNSUPPORTED, "head_dim %d must be a multiple of 8", head_dim);
  if (dtype != FI_DTYPE_F16 && dtype != FI_DTYPE_BF16 && dtype != FI_DTYPE_F32)
    return fail(FI_ERR_UNSUPPORTED, "merge ops support f16 / bf16 / f32 values, got dtype %d", dtype);
  return FI_OK;
}

static int grid_for(int64_t work_items, int device) {
  int sms = 148;
  device_sm_count(device, &sms);
  int64_t blocks = (work_items + 255) / 256;
  const int64_t cap = (int64_t)sms * 8;  // grid-stride beyond 8 CTAs/SM
  if (blocks > cap) blocks = cap;
  return blocks < 1 ? 1 : (int)blocks;
}

}  // namespace fib200

using namespace fib200;

#define DISPATCH_DT(dt, DT, ...)                         \
  switch (dt) {                                          \
    case FI_DTYPE_F16: { constexpr int DT = 0; __VA_ARGS__; break; }  \
    case FI_DTYPE_BF16: { constexpr int DT = 1; __VA_ARGS__; break; } \
    default: { constexpr int DT = 4; __VA_ARGS__; break; }            \
  }

extern "C" {

int fi_merge_state(const void* v_a, const float* s_a, const void* v_b, const float* s_b,
                   void* v_merged, float* s_merged, int seq_len, int num_heads, int head_dim,
                   int dtype, int device, void* stream) {
  int rc = check_merge_shape(seq_len, num_heads, head_dim, dtype);
  if (rc != FI_OK) return rc;
  if (seq_len == 0) return FI_OK;
  FI_CUDA_TRY(cudaSetDevice(device));
  const int64_t rows = (int64_t)seq_len * num_heads;
  const int grid = grid_for(rows * (head_dim / 8), device);
  DISPATCH_DT(dtype, DT,
              (merge_state_kernel<DT, false><<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
                  v_a, s_a, v_b, s_b, v_merged, s_merged, nullptr, rows, num_heads, head_dim)));
  FI_CUDA_TRY(cudaGetLastError());
  return FI_OK;
}

int fi_merge_state_in_place(void* v, float* s, const void* v_other, const float* s_other,
                            const uint8_t* mask_or_null, int seq_len, int num_heads, int head_dim,
                            int dtype, int device, void* stream) {
  int rc = check_merge_shape(seq_len, num_heads, head_dim, dtype);
  if (rc != FI_OK) return rc;
  if (seq_len == 0) return FI_OK;
  FI_CUDA_TRY(cudaSetDevice(device));
  const int64_t rows = (int64_t)seq_len * num_heads;
  const int grid = grid_for(rows * (head_dim / 8), device);
  DISPATCH_DT(dtype, DT,
              (merge_state_kernel<DT, true><<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
                  v, s, v_other, s_other, v, s, mask_or_null, rows, num_heads, head_dim)));
  FI_CUDA_TRY(cudaGetLastError());
  return FI_OK;
}

int fi_merge_states(const void* v, const float* s, void* v_merged, float* s_merged, int seq_len,
                    int num_index_sets, int num_heads, int head_dim, int dtype, int device,
                    void* stream) {
  int rc = check_merge_shape(seq_len, num_heads, head_dim, dtype);
  if (rc != FI_OK) return rc;
  if (num_index_sets < 0) return fail(FI_ERR_INVALID_ARGUMENT, "negative num_index_sets");
  if (seq_len == 0) return FI_OK;
  FI_CUDA_TRY(cudaSetDevice(device));
  const int64_t rows = (int64_t)seq_len * num_heads;
  const int grid = grid_for(rows * (head_dim / 8), device);
  DISPATCH_DT(dtype, DT,
              (merge_states_kernel<DT><<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
                  v, s, v_merged, s_merged, seq_len, num_index_sets, num_heads, head_dim)));
  FI_CUDA_TRY(cudaGetLastError());
  return FI_OK;
}

int fi_variable_length_merge_states(const void* v, const float* s, const int32_t* indptr,
                                    void* v_merged, float* s_merged, int seq_len, int num_heads,
                                    int head_dim, int dtype_in, int dtype_out, int skip_empty,
                                    float out_scale, int device, void* stream) {
  int rc = check_merge_shape(seq_len, num_heads, head_dim, dtype_in);
  if (rc != FI_OK) return rc;
  rc = check_merge_shape(seq_len, num_heads, head_dim, dtype_out);
  if (rc != FI_OK) return rc;
  if (seq_len == 0) return FI_OK;
  FI_CUDA_TRY(cudaSetDevice(device));
  const int64_t rows = (int64_t)seq_len * num_heads;
  const int grid = grid_for(rows * (head_dim / 8), device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  DISPATCH_DT(dtype_in, DI, {
    DISPATCH_DT(dtype_out, DO,
                (varlen_merge_states_kernel<DI, DO><<<grid, 256, 0, st>>>(
                    v, s, indptr, v_merged, s_merged, seq_len, num_heads, head_dim, skip_empty,
                    out_scale)));
  });
  FI_CUDA_TRY(cudaGetLastError());
  return FI_OK;
}

}  // extern "C"
