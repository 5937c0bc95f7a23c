// Library-wide plumbing: thread-local error text, device check, tunables.
#include "common.cuh"
#include <string.h>

namespace b200grpo {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

static int g_sm_count = 0;

int sm_count() {
  if (g_sm_count > 0) return g_sm_count;
  int dev = 0, n = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
  g_sm_count = n;
  return n;
}

int check_device_sm100() {
  static thread_local int cached = 1;  // 1 = unknown
  if (cached <= 0) return cached;
  int dev = 0, major = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e == cudaSuccess) e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
  if (e != cudaSuccess) {
    set_error("no usable CUDA device: %s (libb200grpo has no CPU fallback)", cudaGetErrorString(e));
    (void)cudaGetLastError();
    return B200GRPO_ECUDA;
  }
  if (major != 10) {
    set_error("device compute capability %d.x is not sm_100 (libb200grpo is B200-only)", major);
    return B200GRPO_EARCH;
  }
  cached = 0;
  return 0;
}

// fwd_splits: 0 = auto; two_cta: 1 = cta_group::2 pairs; gemm_stages: 0 = auto; watchdog_ms
static int g_tunables[TUNE_COUNT] = {0, 0, 0, 4000};
int get_tunable(int id) { return g_tunables[id]; }

}  // namespace b200grpo

using namespace b200grpo;

extern "C" const char* b200grpo_last_error(void) { return g_err; }
extern "C" int b200grpo_abi_version(void) { return B200GRPO_ABI_VERSION; }

extern "C" int b200grpo_set_tunable(const char* name, int value) {
  static const char* names[TUNE_COUNT] = {"fwd_splits", "two_cta", "gemm_stages", "watchdog_ms"};
  for (int i = 0; i < TUNE_COUNT; ++i)
    if (name && strcmp(name, names[i]) == 0) {
      const int prev = g_tunables[i];
      g_tunables[i] = value;
      return prev;
    }
  set_error("unknown tunable '%s'", name ? name : "(null)");
  return B200GRPO_EINVAL;
}
