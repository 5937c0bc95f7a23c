// Segment boundaries of a packed row, bit-exact with
// /root/reference/pipelinerl/finetune/rl/__init__.py:159-170:
//   starts = where(position_ids == 0) with index 0 forced; bounds = cat(starts, [T]).
// Three small launches (count -> scan -> scatter); ordered output, no atomics.
#include "common.cuh"

namespace b200grpo {

constexpr int SB_THREADS = 256;
constexpr int SB_ITEMS = 4;  // 1024 positions per block

__device__ __forceinline__ bool is_start(const int64_t* pos, int i) { return i == 0 || pos[i] == 0; }

__global__ void __launch_bounds__(SB_THREADS) seg_count_kernel(const int64_t* pos, int T, int* block_counts) {
  const int base = blockIdx.x * SB_THREADS * SB_ITEMS;
  int c = 0;
#pragma unroll
  for (int k = 0; k < SB_ITEMS; ++k) {
    const int i = base + k * SB_THREADS + threadIdx.x;
    if (i < T && is_start(pos, i)) ++c;
  }
  c = (int)warp_sum((float)c);
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

__global__ void __launch_bounds__(SB_THREADS) seg_scatter_kernel(const int64_t* pos, int T, const int* block_offsets,
                                                                 const int32_t* n_bounds, int32_t* bounds) {
  // positions are visited in index order: thread t owns SB_ITEMS consecutive positions
  const int base = blockIdx.x * SB_THREADS * SB_ITEMS + threadIdx.x * SB_ITEMS;
  bool f[SB_ITEMS];
  int c = 0;
#pragma unroll
  for (int k = 0; k < SB_ITEMS; ++k) { f[k] = (base + k < T) && is_start(pos, base + k); c += f[k]; }
  __shared__ int sh[SB_THREADS];
  sh[threadIdx.x] = c;
  __syncthreads();
  for (int o = 1; o < SB_THREADS; o <<= 1) {
    const int add = threadIdx.x >= o ? sh[threadIdx.x - o] : 0;
    __syncthreads();
    sh[threadIdx.x] += add;
    __syncthreads();
  }
  int w = block_offsets[blockIdx.x] + sh[threadIdx.x] - c;
#pragma unroll
  for (int k = 0; k < SB_ITEMS; ++k) if (f[k]) bounds[w++] = base + k;
  if (blockIdx.x == 0 && threadIdx.x == 0) bounds[*n_bounds - 1] = T;
}

}  // namespace b200grpo

using namespace b200grpo;

extern "C" int b200grpo_segment_bounds(const int64_t* position_ids, int T, int32_t* bounds, int32_t* n_bounds,
                                       void* workspace, size_t workspace_bytes, void* stream) {
  if (int e = check_device_sm100()) return e;
  B200_REQUIRE(position_ids && bounds && n_bounds && workspace && T >= 1, "segment_bounds: bad argument");
  const int n_blocks = ceil_div(T, SB_THREADS * SB_ITEMS);
  if (workspace_bytes < (size_t)n_blocks * sizeof(int)) { set_error("segment_bounds: workspace too small"); return B200GRPO_EWORKSPACE; }
  cudaStream_t st = (cudaStream_t)stream;
  int* counts = (int*)workspace;
  ProfScope ps(PROF_SEGMENTS, st);
  seg_count_kernel<<<n_blocks, SB_THREADS, 0, st>>>(position_ids, T, counts);
  seg_scan_kernel<<<1, 1024, 0, st>>>(counts, n_blocks, n_bounds);
  seg_scatter_kernel<<<n_blocks, SB_THREADS, 0, st>>>(position_ids, T, counts, n_bounds, bounds); count_launch(3);
  B200_CUDA_OK(cudaGetLastError());
  return B200GRPO_OK;
}
