// Shared host/device helpers for libb200grpo (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>

#include "../../include/b200grpo.h"

namespace b200grpo {

// ---- error plumbing -----------------------------------------------------------------
void set_error(const char* fmt, ...);
int check_device_sm100();  // 0 or B200GRPO_EARCH / ECUDA (cached per thread)
int get_tunable(int id);
enum { TUNE_TWO_CTA = 0, TUNE_DH_KSPLIT, TUNE_DW_KSPLIT, TUNE_TAIL_SPLIT, TUNE_FWD_GROUP_M, TUNE_DH_GROUP_M, TUNE_DW_GROUP_M,
       TUNE_DW_GROUP_N, TUNE_SAVE_LOGITS, TUNE_HINT_RESIDENT, TUNE_HINT_STREAM, TUNE_DH_BIG_M, TUNE_DW_BIG_M, TUNE_GEMM_BIG_M,
       TUNE_FWD_BIG_M, TUNE_DW_ZIGZAG, TUNE_ROWS_PATH, TUNE_LOCKSTEP_FWD, TUNE_LOCKSTEP_DH, TUNE_LOCKSTEP_DW, TUNE_LOCKSTEP_SLACK, TUNE_COUNT };

#define B200_CUDA_OK(expr)                                                              \
  do {                                                                                  \
    cudaError_t _e = (expr);                                                            \
    if (_e != cudaSuccess) {                                                            \
      ::b200grpo::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e),     \
                            __FILE__, __LINE__);                                        \
      return B200GRPO_ECUDA;                                                            \
    }                                                                                   \
  } while (0)

#define B200_REQUIRE(cond, ...)                                                         \
  do {                                                                                  \
    if (!(cond)) {                                                                      \
      ::b200grpo::set_error(__VA_ARGS__);                                               \
      return B200GRPO_EINVAL;                                                           \
    }                                                                                   \
  } while (0)

static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
int sm_count();  // multiprocessors of the current device (cached)
// One-time set-up that is PER DEVICE (kernel function attributes such as the dynamic shared-memory limit belong to the
// device's context): first() is true the first time it is asked on the current device.  A process that drives several
// GPUs configures each of them; a racing second thread at worst repeats a harmless cudaFuncSetAttribute.
struct PerDeviceOnce {
  unsigned long long done = 0;
  bool first() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return true;
    const unsigned long long bit = 1ull << dev;
    if (done & bit) return false;
    done |= bit;
    return true;
  }
};

void count_launch(int n = 1);  // kernel launches issued by this library (b200grpo_launch_count)

// ---- optional per-launch timing (b200grpo_profile_*): CUDA events on the launching stream ----
enum { PROF_FWD_GEMM = 1, PROF_FWD_MERGE = 2, PROF_FINALIZE = 3, PROF_TOKEN_LOSS = 4, PROF_BWD_DLOGITS = 5,
       PROF_BWD_DH = 6, PROF_BWD_DW = 7, PROF_GEMM = 8, PROF_ADVANTAGE = 9, PROF_BUCKET = 10, PROF_SEGMENTS = 11,
       PROF_FUSED_CHUNK = 12, PROF_DLOGITS_EW = 13, PROF_REDUCE = 14, PROF_INGEST = 15, PROF_FINAL_NORM = 16, PROF_MC_ALLREDUCE = 17 };
struct ProfScope {  // records start/end events around the launches issued while it is alive (no-op when disabled)
  ProfScope(int tag, cudaStream_t st);
  ~ProfScope();
  int idx;
  cudaStream_t st;
};

// ---- device helpers -------------------------------------------------------------------
#ifdef __CUDACC__
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ unsigned warp_or(unsigned v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v |= __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
#endif

}  // namespace b200grpo
