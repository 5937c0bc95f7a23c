// Vocab-parallel exchange over NVSwitch multicast instead of NCCL.
//
// The path (new capability, no reference lines: the reference always holds the whole lm_head -- SURVEY 2.1) has two
// exchange steps per micro-batch:
//   forward : every rank's per-token partials (max, sum-exp, sum e*z, z_target) -> all ranks.  Done INSIDE the merge
//             kernel of K1 (fwd_merge_kernel<true>, lmhead.cu): the merged row is stored with multimem.st into this
//             rank's slot of a symmetric [R][M][4] buffer, so the all-gather is the kernel's own store.
//   backward: dH = sum over vocab shards.  The dH GEMM writes its partial straight into a symmetric bf16 [M, H] buffer;
//             the kernel below is the two-shot in-switch all-reduce: rank r owns 1/R of the 16-byte vectors,
//             multimem.ld_reduce pulls their f32-accumulated sum over all ranks' copies through the switch,
//             multimem.st writes it back into every copy.  No shared memory and <= 32 registers, so its CTAs co-reside
//             with the dW GEMM's persistent CTAs (one per SM, ~200 KB of shared memory each) instead of taking SMs
//             away from it, which is what an NCCL all-reduce running beside that GEMM does.
// Cross-rank ordering (all partials written before anyone reduces; all slices written back before anyone reads) is the
// caller's symmetric-memory barrier on the same stream.
#include "common.cuh"

namespace b200grpo {

constexpr int MC_THREADS = 512;
constexpr int MC_UNROLL = 4;

__device__ __forceinline__ uint4 multimem_ld_reduce_bf16x8(const uint4* mc) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st_bf16x8(uint4* mc, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1, %2, %3, %4};" ::"l"(mc), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

__global__ void __launch_bounds__(MC_THREADS) mc_allreduce_bf16_kernel(uint4* __restrict__ mc, long long v_begin, long long v_end) {
  const long long stride = (long long)gridDim.x * MC_THREADS;
  for (long long i = v_begin + (long long)blockIdx.x * MC_THREADS + threadIdx.x; i < v_end; i += stride * MC_UNROLL) {
    uint4 v[MC_UNROLL];
#pragma unroll
    for (int u = 0; u < MC_UNROLL; ++u)
      if (i + u * stride < v_end) v[u] = multimem_ld_reduce_bf16x8(mc + i + u * stride);
#pragma unroll
    for (int u = 0; u < MC_UNROLL; ++u)
      if (i + u * stride < v_end) multimem_st_bf16x8(mc + i + u * stride, v[u]);
  }
}

}  // namespace b200grpo

using namespace b200grpo;

extern "C" int b200grpo_mc_allreduce_bf16(void* mc_ptr, int64_t numel, int rank, int world, int ctas, void* stream) {
  if (int e = check_device_sm100()) return e;
  B200_REQUIRE(mc_ptr && ((uintptr_t)mc_ptr & 15) == 0 && numel >= 8 && numel % 8 == 0, "mc_allreduce: 16-byte aligned multicast pointer, numel %% 8 == 0");
  B200_REQUIRE(world >= 1 && rank >= 0 && rank < world, "mc_allreduce: rank %d of %d", rank, world);
  const long long nvec = numel / 8;
  const long long per = (nvec + world - 1) / world;
  long long b = per * rank, e = b + per;
  if (b > nvec) b = nvec;
  if (e > nvec) e = nvec;
  if (ctas <= 0) ctas = 16;
  const long long want = (e - b + (long long)MC_THREADS * MC_UNROLL - 1) / ((long long)MC_THREADS * MC_UNROLL);
  if (want < ctas) ctas = (int)(want < 1 ? 1 : want);
  ProfScope ps(PROF_MC_ALLREDUCE, (cudaStream_t)stream);
  mc_allreduce_bf16_kernel<<<ctas, MC_THREADS, 0, (cudaStream_t)stream>>>((uint4*)mc_ptr, b, e); count_launch();
  B200_CUDA_OK(cudaGetLastError());
  return B200GRPO_OK;
}
