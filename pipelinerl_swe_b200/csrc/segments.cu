// Segment boundaries of a packed row, bit-exact with
// /root/reference/pipelinerl/finetune/rl/__init__.py:159-170:
//   starts = where(position_ids == 0) with index 0 forced; bounds = cat(starts, [T]).
// Three small launches (count -> scan -> scatter); ordered output, no atomics.  position_ids is read ONCE (8 B/token,
// eight independent coalesced loads in flight per thread): the count pass also leaves a 1-bit-per-token start mask
// (one ballot word per warp and trip, 1/8 B per token) and the scatter pass works from that mask alone.
#include "common.cuh"
#include <mutex>

namespace b200grpo {

constexpr int SB_THREADS = 256;
constexpr int SB_ITEMS = 8;                         // 2048 positions per block
constexpr int SB_WORDS = SB_THREADS * SB_ITEMS / 32;  // 64 mask words per block

__global__ void __launch_bounds__(SB_THREADS) seg_count_kernel(const int64_t* __restrict__ pos, int T, int* __restrict__ block_counts,
                                                               uint32_t* __restrict__ mask) {
  const int base = blockIdx.x * SB_THREADS * SB_ITEMS;
  int64_t v[SB_ITEMS];
#pragma unroll
  for (int k = 0; k < SB_ITEMS; ++k) {   // all loads issued before any is used
    const int i = base + k * SB_THREADS + threadIdx.x;
    v[k] = i < T ? __ldg(pos + i) : 1;
  }
  int c = 0;
#pragma unroll
  for (int k = 0; k < SB_ITEMS; ++k) {
    const int i = base + k * SB_THREADS + threadIdx.x;
    const bool f = i < T && (i == 0 || v[k] == 0);
    const unsigned m = __ballot_sync(0xffffffffu, f);
    if ((threadIdx.x & 31) == 0) {
      mask[(base + k * SB_THREADS + (int)threadIdx.x) >> 5] = m;   // word of tokens [i, i + 32)
      c += __popc(m);
    }
  }
  __shared__ int sh[SB_THREADS / 32];
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int w = 0; w < SB_THREADS / 32; ++w) t += sh[w];
    block_counts[blockIdx.x] = t;
  }
}

// single block: exclusive scan of block_counts in place; writes S+1 to n_bounds
__global__ void __launch_bounds__(1024) seg_scan_kernel(int* block_counts, int n_blocks, int32_t* n_bounds) {
  __shared__ int sh[1024];
  __shared__ int carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int base = 0; base < n_blocks; base += 1024) {
    const int i = base + threadIdx.x;
    const int v = i < n_blocks ? block_counts[i] : 0;
    sh[threadIdx.x] = v;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {  // Hillis-Steele inclusive scan
      const int add = threadIdx.x >= o ? sh[threadIdx.x - o] : 0;
      __syncthreads();
      sh[threadIdx.x] += add;
      __syncthreads();
    }
    if (i < n_blocks) block_counts[i] = carry + sh[threadIdx.x] - v;
    __syncthreads();
    if (threadIdx.x == 1023) carry += sh[1023];
    __syncthreads();
  }
  if (threadIdx.x == 0) *n_bounds = carry + 1;
}

// one thread per mask word (32 positions), SB_WORDS words per block == one count block: ordered emission of the set bits
__global__ void __launch_bounds__(SB_WORDS) seg_scatter_kernel(const uint32_t* __restrict__ mask, int T, const int* __restrict__ block_offsets,
                                                               const int32_t* __restrict__ n_bounds, int32_t* __restrict__ bounds) {
  const int word = blockIdx.x * SB_WORDS + threadIdx.x;
  uint32_t m = (word * 32 < T) ? mask[word] : 0u;
  const int c = __popc(m);
  // exclusive scan of the SB_WORDS (= 64) popcounts: two warps
  int incl = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int up = __shfl_up_sync(0xffffffffu, incl, o);
    if ((threadIdx.x & 31) >= o) incl += up;
  }
  __shared__ int warp_total[SB_WORDS / 32];
  if ((threadIdx.x & 31) == 31) warp_total[threadIdx.x >> 5] = incl;
  __syncthreads();
  int w = block_offsets[blockIdx.x] + incl - c;
  for (int k = 0; k < (int)(threadIdx.x >> 5); ++k) w += warp_total[k];
  while (m) {
    bounds[w++] = word * 32 + __ffs(m) - 1;
    m &= m - 1;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) bounds[*n_bounds - 1] = T;
}

}  // namespace b200grpo

using namespace b200grpo;

static size_t seg_counts_bytes(int T) { return align_up((size_t)ceil_div(T, SB_THREADS * SB_ITEMS) * sizeof(int), 256); }

extern "C" size_t b200grpo_segment_bounds_workspace(int T) {
  if (T < 1) T = 1;
  return seg_counts_bytes(T) + align_up((size_t)ceil_div(T, SB_THREADS * SB_ITEMS) * SB_WORDS * sizeof(uint32_t), 256);
}

extern "C" int b200grpo_segment_bounds(const int64_t* position_ids, int T, int32_t* bounds, int32_t* n_bounds,
                                       void* workspace, size_t workspace_bytes, void* stream) {
  if (int e = check_device_sm100()) return e;
  B200_REQUIRE(position_ids && bounds && n_bounds && workspace && T >= 1, "segment_bounds: bad argument");
  const int n_blocks = ceil_div(T, SB_THREADS * SB_ITEMS);
  if (workspace_bytes < b200grpo_segment_bounds_workspace(T)) { set_error("segment_bounds: workspace too small"); return B200GRPO_EWORKSPACE; }
  cudaStream_t st = (cudaStream_t)stream;
  int* counts = (int*)workspace;
  uint32_t* mask = (uint32_t*)((char*)workspace + seg_counts_bytes(T));
  ProfScope ps(PROF_SEGMENTS, st);
  seg_count_kernel<<<n_blocks, SB_THREADS, 0, st>>>(position_ids, T, counts, mask);
  seg_scan_kernel<<<1, 1024, 0, st>>>(counts, n_blocks, n_bounds);
  seg_scatter_kernel<<<n_blocks, SB_WORDS, 0, st>>>(mask, T, counts, n_bounds, bounds); count_launch(3);
  B200_CUDA_OK(cudaGetLastError());
  return B200GRPO_OK;
}

// ---- allocation-free variant -------------------------------------------------------------------------------------------
// No counterpart in the reference: added for callers that cannot call cudaMalloc
// (CUDA graph capture, tight serving loops, harnesses that forbid it).  Same three kernels, scratch (per-block counts +
// the 1-bit start mask) in a file-scope __device__ array sized for B200GRPO_SEGMENT_BOUNDS_STATIC_MAX_T tokens; every GPU
// that loads the library holds those ~520 KB until exit.  In host code the array's NAME is a host-side stand-in, not a
// device address: its address on the current device comes from cudaGetSymbolAddress, once per device, below.
// Every call shares the one array: calls that overlap on the GPU (two streams, two threads) race.  Nothing in it has to
// start at zero -- each count and each mask word that a call reads is written by that call's own count pass first.
namespace b200grpo {
constexpr int SEG_STATIC_BLOCKS = B200GRPO_SEGMENT_BOUNDS_STATIC_MAX_T / (SB_THREADS * SB_ITEMS);
static_assert(B200GRPO_SEGMENT_BOUNDS_STATIC_MAX_T % (SB_THREADS * SB_ITEMS) == 0, "static scratch: whole blocks");
constexpr size_t SEG_STATIC_COUNTS_BYTES = ((size_t)SEG_STATIC_BLOCKS * sizeof(int) + 255) / 256 * 256;
constexpr size_t SEG_STATIC_MASK_BYTES = (size_t)SEG_STATIC_BLOCKS * SB_WORDS * sizeof(uint32_t);
static __device__ __align__(256) unsigned char g_seg_static_scratch[SEG_STATIC_COUNTS_BYTES + SEG_STATIC_MASK_BYTES];

constexpr int SEG_STATIC_MAX_DEVICES = 64;
static std::mutex g_seg_static_mu;
static void* g_seg_static_addr[SEG_STATIC_MAX_DEVICES] = {};

// device address of the scratch array on the CURRENT device (resolved once per device, with that device current)
static int seg_static_scratch(void** out) {
  int dev = -1;
  B200_CUDA_OK(cudaGetDevice(&dev));
  B200_REQUIRE(dev >= 0 && dev < SEG_STATIC_MAX_DEVICES, "segment_bounds_static: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lk(g_seg_static_mu);
  if (!g_seg_static_addr[dev]) {
    void* p = nullptr;
    B200_CUDA_OK(cudaGetSymbolAddress(&p, g_seg_static_scratch));
    B200_REQUIRE(p != nullptr, "segment_bounds_static: cudaGetSymbolAddress returned a null address");
    g_seg_static_addr[dev] = p;
  }
  *out = g_seg_static_addr[dev];
  return B200GRPO_OK;
}
}  // namespace b200grpo

extern "C" int b200grpo_segment_bounds_static(const int64_t* position_ids, int T, int32_t* bounds, int32_t* n_bounds,
                                              void* stream) {
  if (int e = check_device_sm100()) return e;
  B200_REQUIRE(position_ids && bounds && n_bounds && T >= 1, "segment_bounds_static: bad argument");
  if (T > B200GRPO_SEGMENT_BOUNDS_STATIC_MAX_T) {
    set_error("segment_bounds_static: T = %d exceeds the built-in scratch (at most %d tokens); use b200grpo_segment_bounds "
              "with a caller-provided workspace", T, B200GRPO_SEGMENT_BOUNDS_STATIC_MAX_T);
    return B200GRPO_EWORKSPACE;
  }
  void* scratch = nullptr;
  if (int e = seg_static_scratch(&scratch)) return e;
  const int n_blocks = ceil_div(T, SB_THREADS * SB_ITEMS);
  cudaStream_t st = (cudaStream_t)stream;
  int* counts = (int*)scratch;
  uint32_t* mask = (uint32_t*)((char*)scratch + SEG_STATIC_COUNTS_BYTES);
  seg_count_kernel<<<n_blocks, SB_THREADS, 0, st>>>(position_ids, T, counts, mask);
  seg_scan_kernel<<<1, 1024, 0, st>>>(counts, n_blocks, n_bounds);
  seg_scatter_kernel<<<n_blocks, SB_WORDS, 0, st>>>(mask, T, counts, n_bounds, bounds); count_launch(3);
  B200_CUDA_OK(cudaGetLastError());
  return B200GRPO_OK;
}
