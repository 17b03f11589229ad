// host_common.h — host-side helpers shared by the C-ABI translation units:
// thread-local error string, CUDA error mapping, device attribute cache and the
// kernel configuration (stages / CTAs per SM) both plan() and run() must agree on.
#pragma once
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/fi_b200.h"
#include "decode_params.h"

namespace fib200 {

void set_error(const char* fmt, ...);
int fail(int code, const char* fmt, ...);

#define FI_CUDA_TRY(expr)                                                        \
  do {                                                                           \
    cudaError_t _e = (expr);                                                     \
    if (_e != cudaSuccess)                                                       \
      return ::fib200::fail(FI_ERR_CUDA, "%s failed: %s (%s:%d)", #expr,         \
                            cudaGetErrorString(_e), __FILE__, __LINE__);         \
  } while (0)

inline int dtype_size(int dt) {
  switch (dt) {
    case FI_DTYPE_F16:
    case FI_DTYPE_BF16: return 2;
    case FI_DTYPE_E4M3:
    case FI_DTYPE_E5M2: return 1;
    case FI_DTYPE_F32: return 4;
    default: return 0;
  }
}
// KVF template code: 0 = same 16-bit type as q, 1 = e4m3, 2 = e5m2, -1 = unsupported pair
inline int kv_format(int q_dtype, int kv_dtype) {
  if (kv_dtype == q_dtype && (q_dtype == FI_DTYPE_F16 || q_dtype == FI_DTYPE_BF16)) return 0;
  if (q_dtype != FI_DTYPE_F16 && q_dtype != FI_DTYPE_BF16) return -1;
  if (kv_dtype == FI_DTYPE_E4M3) return 1;
  if (kv_dtype == FI_DTYPE_E5M2) return 2;
  return -1;
}

struct KernelConfig {
  int stages;
  int ctas_per_sm;
  int smem_bytes;
};
// Shared-memory budget: 227 KiB usable per SM, 1 KiB reserved per resident CTA.
KernelConfig choose_config(int kvf, int head_dim);
int device_sm_count(int device, int* out);

// process-wide tuning overrides (0 = automatic); see fi_set_tuning
struct Tuning {
  int stages;
  int ctas_per_sm;
  int loader;          // 0 auto, 1 force TMA, 2 force cp.async
  int min_tiles_per_cta;
};
Tuning& tuning();

// Item scheduler knobs (environment, read once): FI_B200_SCHED_F0 / _G (percent),
// FI_B200_SCHED_TAIL / _LOOKAHEAD (tiles).  See plan.cu.
struct SchedTuning {
  int f0_pct;           // share of a CTA's fair share handed out statically (100: static split)
  int g_pct;            // share of the remainder cut per later round
  int tail_tiles;       // smallest item of the later rounds
  int lookahead_tiles;  // kernel: draw the next item when this many published tiles are left
};
const SchedTuning& sched_tuning();

}  // namespace fib200
