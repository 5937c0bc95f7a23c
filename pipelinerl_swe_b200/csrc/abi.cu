// Library-wide plumbing: thread-local error text, device check, tunables.
#include "common.cuh"
#include <string.h>
#include <vector>

namespace b200grpo {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

static unsigned long long g_launches = 0;  // relaxed counter; exactness under concurrent host threads is not needed
void count_launch(int n) { __atomic_fetch_add(&g_launches, (unsigned long long)n, __ATOMIC_RELAXED); }
unsigned long long launches_so_far() { return __atomic_load_n(&g_launches, __ATOMIC_RELAXED); }

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

// Tunables, in the order of the TUNE_* enum (common.cuh); defaults are what ships.
//   two_cta: 1 = cta_group::2 CTA pairs;  dh_ksplit / dw_ksplit: 0 = automatic (tail split when tail_split = 1, else the
//   uniform split count with the best wave efficiency), >= 1 = that many uniform k-splits;
//   fwd_group_m / dh_group_m / dw_group_m: row blocks per raster group; dw_group_n: column blocks per group (0 = all);
//   save_logits: used by b200grpo_rl_step_host; hint_resident / hint_stream: L2 eviction policy of reused / streamed
//   operands (0 normal, 1 evict-first, 2 evict-last);
//   dh_big_m / dw_big_m / gemm_big_m / fwd_big_m: 256 rows per CTA (pair tile 512 x 256, one accumulator stage).  dH ships with it
//   (long K loop, plain store epilogue); dW does not: its f32 accumulate epilogue must overlap the next mainloop.
//   dw_zigzag: odd tiles of a cluster walk the token (K) axis of dW backwards, so a wave starts on the rows of `hidden`
//   the previous wave touched last (L2 reuse instead of a DRAM re-read).  Ships on: dW DRAM reads 12.0 -> 10.4 GB, dW -1 %
//   same-box (profiles/r01/dw_zigzag_ab.md).
//   rows_path: an entropy-free backward (coef_ent == NULL) runs its GEMMs on the saved exponentials, the onehot terms added after
//   them (dH epilogue gather + sparse target-row kernel): 2 = except for confident targets (p > 1/2), whose onehot is folded into
//   the saved exponential before the GEMMs (shipped), 1 = for every row; 0 = dlogits formed in place first like the entropy
//   path (A/B switches).
//   lockstep_fwd / lockstep_dh / lockstep_dw: k-blocks per epoch of the soft K-lockstep between the clusters of a wave
//   (gemm_sm100.cuh lockstep_pace; 0 = off) for the forward / recompute GEMM, dH and dW; lockstep_slack: epochs a cluster may
//   run ahead of the slowest one.
static const char* const g_tunable_names[TUNE_COUNT] = {"two_cta", "dh_ksplit", "dw_ksplit", "tail_split", "fwd_group_m", "dh_group_m",
                                                        "dw_group_m", "dw_group_n", "save_logits", "hint_resident", "hint_stream",
                                                        "dh_big_m", "dw_big_m", "gemm_big_m", "fwd_big_m", "dw_zigzag", "rows_path", "lockstep_fwd",
                                                        "lockstep_dh", "lockstep_dw", "lockstep_slack"};
static int g_tunables[TUNE_COUNT] = {1, 0, 0, 1, 16, 4, 1, 0, 1, 0, 0, 1, 0, 0, 0, 1, 2, 16, 8, 32, 2};
int get_tunable(int id) { return g_tunables[id]; }

// ---- profiling ring (diagnostic; single-threaded use) -------------------------------------
struct ProfSlot { cudaEvent_t a, b; int tag; };
static std::vector<ProfSlot> g_prof;
static int g_prof_used = 0;
static bool g_prof_on = false;

ProfScope::ProfScope(int tag, cudaStream_t s) : idx(-1), st(s) {
  if (!g_prof_on || g_prof_used >= (int)g_prof.size()) return;
  idx = g_prof_used++;
  g_prof[idx].tag = tag;
  cudaEventRecord(g_prof[idx].a, st);
}
ProfScope::~ProfScope() {
  if (idx >= 0) cudaEventRecord(g_prof[idx].b, st);
}

}  // namespace b200grpo

using namespace b200grpo;

extern "C" int b200grpo_profile_enable(int capacity) {
  if (int e = check_device_sm100()) return e;
  g_prof_on = capacity > 0;
  g_prof_used = 0;
  while ((int)g_prof.size() < capacity) {
    ProfSlot s;
    B200_CUDA_OK(cudaEventCreate(&s.a));
    B200_CUDA_OK(cudaEventCreate(&s.b));
    s.tag = 0;
    g_prof.push_back(s);
  }
  return B200GRPO_OK;
}

extern "C" int b200grpo_profile_read(int32_t* tags_host, float* ms_host, int capacity) {
  const int n = g_prof_used < capacity ? g_prof_used : capacity;
  for (int i = 0; i < n; ++i) {
    B200_CUDA_OK(cudaEventSynchronize(g_prof[i].b));
    float ms = 0.f;
    B200_CUDA_OK(cudaEventElapsedTime(&ms, g_prof[i].a, g_prof[i].b));
    tags_host[i] = g_prof[i].tag;
    ms_host[i] = ms;
  }
  g_prof_used = 0;
  return n;
}

extern "C" const char* b200grpo_last_error(void) { return g_err; }
extern "C" int b200grpo_abi_version(void) { return B200GRPO_ABI_VERSION; }

extern "C" uint64_t b200grpo_launch_count(void) { return b200grpo::launches_so_far(); }

extern "C" int b200grpo_set_tunable(const char* name, int value) {
  const char* const* names = g_tunable_names;
  for (int i = 0; i < TUNE_COUNT; ++i)
    if (name && strcmp(name, names[i]) == 0) {
      const int prev = g_tunables[i];
      g_tunables[i] = value;
      return prev;
    }
  set_error("unknown tunable '%s'", name ? name : "(null)");
  return B200GRPO_EINVAL;
}
