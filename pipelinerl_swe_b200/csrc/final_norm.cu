// Final RMSNorm and the value head in ONE sweep over the last hidden state.
//
// What it replaces on the trainer path: inside `model(**model_inputs)` of
// /root/reference/pipelinerl/finetune/rl/__init__.py:183-195 the trunk's last op is the final RMSNorm
// (third-party: transformers `Qwen2RMSNorm` / `LlamaRMSNorm`, ~7 elementwise launches over [T, H] in eager),
// and /root/reference/pipelinerl/finetune/value_model.py:40-52,100-106 then runs `Linear(H, 1)` over the same
// normalised tensor.  Here: one launch reads the pre-norm hidden state once and writes the normalised bf16
// rows (operand A of K1) plus one value per token; the backward is one launch + a small column reduction.
//
// Numerics follow the transformers module on bf16 inputs: variance and x * rsqrt(var + eps) in f32, the
// product rounded to bf16, then `weight * x` as a bf16 multiply (f32 product rounded to bf16).
//
// Why this is a separate pre-pass and NOT a prologue of the K1 GEMM: K1 re-loads every A tile once per vocab
// tile (594 times at V = 152064), so normalising inside the operand pipeline would redo the transform 594x
// and put a shared-memory round trip between TMA and tcgen05.mma; the per-row scale could move to the
// epilogue but the per-column gain cannot (it multiplies along K).  Normalising once costs 4 B/element.
//
// HBM-bound byte work.  Forward: 4 B/element (bf16 in + out).  Backward: 6 B/element (x, dy in; dx out).
#include "common.cuh"

namespace b200grpo {

constexpr int FN_THREADS = 256;
constexpr int FN_DEPTH = 4;   // rows of (x, dy) in flight per block in the backward's cp.async ring

__device__ __forceinline__ void cp_async_16(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}

__device__ __forceinline__ void unpack8(const uint4& q, float (&f)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&q);
#pragma unroll
  for (int j = 0; j < 4; ++j) { const float2 v = __bfloat1622float2(h[j]); f[2 * j] = v.x; f[2 * j + 1] = v.y; }
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  uint4 q;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&q);
#pragma unroll
  for (int j = 0; j < 4; ++j) h[j] = __floats2bfloat162_rn(f[2 * j], f[2 * j + 1]);
  return q;
}
__device__ __forceinline__ float bf16_round(float v) { return __bfloat162float(__float2bfloat16_rn(v)); }

// ---- forward: one warp per row, streamed twice (the second read of the 7 KB row hits L1/L2).  Holding the row in
// registers instead was measured slower: 185 registers -> 8 warps per SM, 0.50 vs 0.70 of the HBM peak at H = 3584 ----
template <bool kValue>
__global__ void __launch_bounds__(FN_THREADS) final_norm_fwd_kernel(const __nv_bfloat16* __restrict__ x, long long ldx,
                                                                    const __nv_bfloat16* __restrict__ gamma, float eps,
                                                                    const __nv_bfloat16* __restrict__ vw,
                                                                    const __nv_bfloat16* __restrict__ vb, int T, int H,
                                                                    __nv_bfloat16* __restrict__ y, long long ldy,
                                                                    float* __restrict__ rstd, float* __restrict__ value) {
  const int lane = threadIdx.x & 31;
  const int warps = FN_THREADS / 32;
  const int nvec = H >> 3;
  const uint4* g4 = reinterpret_cast<const uint4*>(gamma);
  const uint4* w4 = reinterpret_cast<const uint4*>(vw);
  const float bias = (kValue && vb) ? __bfloat162float(*vb) : 0.f;
  for (int row = blockIdx.x * warps + (threadIdx.x >> 5); row < T; row += gridDim.x * warps) {
    const uint4* xr = reinterpret_cast<const uint4*>(x + (long long)row * ldx);
    uint4* yr = reinterpret_cast<uint4*>(y + (long long)row * ldy);
    float ss = 0.f;
    {
      for (int v0 = lane; v0 < nvec; v0 += 128) {  // four independent 16-byte loads in flight per lane
        uint4 q[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) q[u] = v0 + 32 * u < nvec ? __ldg(xr + v0 + 32 * u) : make_uint4(0, 0, 0, 0);
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          float f[8];
          unpack8(q[u], f);
#pragma unroll
          for (int j = 0; j < 8; ++j) ss = fmaf(f[j], f[j], ss);
        }
      }
    }
    ss = warp_sum(ss);
    const float r = rsqrtf(ss / (float)H + eps);
    float acc = 0.f;
    auto emit = [&](int v, const uint4& q) {
      float f[8], g[8], o[8];
      unpack8(q, f);
      unpack8(__ldg(g4 + v), g);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = bf16_round(g[j] * bf16_round(f[j] * r));
      if constexpr (kValue) {
        float w[8];
        unpack8(__ldg(w4 + v), w);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc = fmaf(o[j], w[j], acc);
      }
      yr[v] = pack8(o);
    };
    for (int v = lane; v < nvec; v += 32) emit(v, xr[v]);  // second read of a row this warp just streamed: L1/L2
    if constexpr (kValue) acc = warp_sum(acc);
    if (lane == 0) {
      rstd[row] = r;
      if constexpr (kValue) value[row] = acc + bias;
    }
  }
}

// ---- backward: a block owns rows blockIdx.x, +grid, ...; thread t owns columns (k*256 + t)*8 .. +8 -----------
// so the column sums (dgamma, dvalue_w) stay in registers across rows; the per-row projection term is one block
// reduction per row (double-buffered: one __syncthreads per row); rows are prefetched FN_DEPTH deep with cp.async.
template <int NCH, int kThreads, bool kValue>
__global__ void __launch_bounds__(kThreads, kThreads == 256 ? 2 : 1)
final_norm_bwd_kernel(const __nv_bfloat16* __restrict__ x, long long ldx, const __nv_bfloat16* __restrict__ gamma,
                      const float* __restrict__ rstd, const __nv_bfloat16* __restrict__ dy, long long lddy,
                      const float* __restrict__ dvalue, const __nv_bfloat16* __restrict__ vw, int T, int H,
                      __nv_bfloat16* __restrict__ dx, long long lddx, float* __restrict__ part, float* __restrict__ part_b) {
  __shared__ float red[2][kThreads / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nvec = H >> 3;
  uint4 g_pk[NCH], w_pk[NCH];
  float dg[NCH][8], dw[NCH][8];
#pragma unroll
  for (int k = 0; k < NCH; ++k) {
    const int v = k * kThreads + tid;
    g_pk[k] = v < nvec ? __ldg(reinterpret_cast<const uint4*>(gamma) + v) : make_uint4(0, 0, 0, 0);
    w_pk[k] = (kValue && v < nvec) ? __ldg(reinterpret_cast<const uint4*>(vw) + v) : make_uint4(0, 0, 0, 0);
#pragma unroll
    for (int j = 0; j < 8; ++j) { dg[k][j] = 0.f; dw[k][j] = 0.f; }
  }
  float db = 0.f;
  int parity = 0;
  // Rows arrive through a cp.async ring in shared memory (FN_DEPTH rows of x and dy per block): every thread copies and
  // later reads only ITS OWN 16-byte chunks, so the ring needs no block barrier and the loads in flight cost no
  // registers.  Operands stay packed (bf16) across the block reduction; the f32 terms are recomputed after it.
  extern __shared__ uint4 ring[];                    // [FN_DEPTH][2][nvec]
  uint4 xq[NCH], dq[NCH];
  auto issue_row = [&](int row, int slot) {
    if (row < T) {
      const uint4* xr = reinterpret_cast<const uint4*>(x + (long long)row * ldx);
      const uint4* dr = reinterpret_cast<const uint4*>(dy + (long long)row * lddy);
      uint4* sx = ring + (size_t)slot * 2 * nvec;
#pragma unroll
      for (int k = 0; k < NCH; ++k) {
        const int v = k * kThreads + tid;
        if (v < nvec) { cp_async_16(sx + v, xr + v); cp_async_16(sx + nvec + v, dr + v); }
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
#pragma unroll
  for (int d = 0; d < FN_DEPTH; ++d) issue_row(blockIdx.x + d * gridDim.x, d);
  int slot = 0;
  for (int row = blockIdx.x; row < T; row += gridDim.x, parity ^= 1) {
    asm volatile("cp.async.wait_group %0;" ::"n"(FN_DEPTH - 1) : "memory");
    {
      const uint4* sx = ring + (size_t)slot * 2 * nvec;
#pragma unroll
      for (int k = 0; k < NCH; ++k) {
        const int v = k * kThreads + tid;
        xq[k] = v < nvec ? sx[v] : make_uint4(0, 0, 0, 0);
        dq[k] = v < nvec ? sx[nvec + v] : make_uint4(0, 0, 0, 0);
      }
    }
    issue_row(row + FN_DEPTH * gridDim.x, slot);     // refill the slot just drained (own chunks only: no hazard)
    slot = slot + 1 == FN_DEPTH ? 0 : slot + 1;
    const float r = __ldg(rstd + row);
    const float dv = kValue ? __ldg(dvalue + row) : 0.f;
    if (kValue && tid == 0) db += dv;
    float dot = 0.f;
#pragma unroll
    for (int k = 0; k < NCH; ++k) {
      float xv[8], dyv[8], g[8], w[8];
      unpack8(xq[k], xv); unpack8(dq[k], dyv); unpack8(g_pk[k], g);
      if constexpr (kValue) unpack8(w_pk[k], w);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float xn = xv[j] * r;            // f32 normalised value (the dx formula)
        const float xnb = bf16_round(xn);      // what the forward multiplied by gamma (the saved tensor in torch)
        float dY = dyv[j];
        if constexpr (kValue) {
          dY = fmaf(dv, w[j], dY);
          dw[k][j] = fmaf(dv, bf16_round(g[j] * xnb), dw[k][j]);
        }
        dg[k][j] = fmaf(dY, xnb, dg[k][j]);
        dot = fmaf(dY * g[j], xn, dot);
      }
    }
    dot = warp_sum(dot);
    if (lane == 0) red[parity][warp] = dot;
    __syncthreads();
    float tot = 0.f;
#pragma unroll
    for (int w = 0; w < kThreads / 32; ++w) tot += red[parity][w];
    const float m = tot / (float)H;
    uint4* out = reinterpret_cast<uint4*>(dx + (long long)row * lddx);
#pragma unroll
    for (int k = 0; k < NCH; ++k) {
      const int v = k * kThreads + tid;
      float xv[8], dyv[8], g[8], w[8], o[8];
      unpack8(xq[k], xv); unpack8(dq[k], dyv); unpack8(g_pk[k], g);
      if constexpr (kValue) unpack8(w_pk[k], w);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float dY = dyv[j];
        if constexpr (kValue) dY = fmaf(dv, w[j], dY);
        o[j] = r * (dY * g[j] - xv[j] * r * m);
      }
      if (v < nvec) out[v] = pack8(o);
    }
  }
  // per-block column partials: part[blockIdx][0][H] = dgamma, [1][H] = dvalue_w
  float* pg = part + (size_t)blockIdx.x * 2 * H;
#pragma unroll
  for (int k = 0; k < NCH; ++k) {
    const int v = k * kThreads + tid;
    if (v < nvec) {
      float4* a = reinterpret_cast<float4*>(pg + v * 8);
      a[0] = make_float4(dg[k][0], dg[k][1], dg[k][2], dg[k][3]);
      a[1] = make_float4(dg[k][4], dg[k][5], dg[k][6], dg[k][7]);
      if constexpr (kValue) {
        float4* b = reinterpret_cast<float4*>(pg + H + v * 8);
        b[0] = make_float4(dw[k][0], dw[k][1], dw[k][2], dw[k][3]);
        b[1] = make_float4(dw[k][4], dw[k][5], dw[k][6], dw[k][7]);
      }
    }
  }
  if (kValue && tid == 0) part_b[blockIdx.x] = db;
}

// fixed-order sum of the block partials (deterministic): one thread per column
__global__ void __launch_bounds__(FN_THREADS) final_norm_reduce_kernel(const float* __restrict__ part, const float* __restrict__ part_b,
                                                                       int n_part, int H, float* __restrict__ dgamma,
                                                                       float* __restrict__ dvw, float* __restrict__ dvb) {
  const int c = blockIdx.x * FN_THREADS + threadIdx.x;
  if (c < H) {
    float a = 0.f, b = 0.f;
    for (int p = 0; p < n_part; ++p) {
      a += part[(size_t)p * 2 * H + c];
      if (dvw) b += part[(size_t)p * 2 * H + H + c];
    }
    dgamma[c] = a;
    if (dvw) dvw[c] = b;
  }
  if (dvb && c == 0) {
    float s = 0.f;
    for (int p = 0; p < n_part; ++p) s += part_b[p];
    *dvb = s;
  }
}

static int fn_bwd_grid(int T, int H) {
  const int per_sm = H <= 4096 ? 2 : 1;   // 256-thread blocks up to H = 4096, one 512-thread block per SM above
  int sms = sm_count();
  if (sms <= 0) sms = 148;
  const int g = sms * per_sm;
  return T < g ? T : g;
}

}  // namespace b200grpo

using namespace b200grpo;

extern "C" size_t b200grpo_final_norm_bwd_workspace(int T, int H) {
  (void)T;
  int sms = sm_count();
  if (sms <= 0) sms = 148;
  return align_up((size_t)sms * 2 * 2 * (size_t)H * sizeof(float), 256) + align_up((size_t)sms * 2 * sizeof(float), 256);
}

static int fn_check(const void* x, int64_t ldx, const void* y, int64_t ldy, int T, int H) {
  B200_REQUIRE(x && y && T >= 1 && H >= 8 && H % 8 == 0 && H <= 8192, "final_norm: T=%d H=%d (H a multiple of 8, <= 8192)", T, H);
  B200_REQUIRE(ldx >= H && ldy >= H && ldx % 8 == 0 && ldy % 8 == 0, "final_norm: leading dimensions must be multiples of 8 elements");
  B200_REQUIRE((((uintptr_t)x | (uintptr_t)y) & 15) == 0, "final_norm: tensors must be 16-byte aligned");
  return B200GRPO_OK;
}

extern "C" int b200grpo_final_norm_fwd(const void* x, int64_t ldx, const void* gamma, float eps, const void* value_w,
                                       const void* value_b, int T, int H, void* y, int64_t ldy, float* rstd, float* value,
                                       void* stream) {
  if (int e = check_device_sm100()) return e;
  if (int e = fn_check(x, ldx, y, ldy, T, H)) return e;
  B200_REQUIRE(gamma && rstd && ((uintptr_t)gamma & 15) == 0, "final_norm_fwd: gamma / rstd");
  B200_REQUIRE(!value_w || (value && ((uintptr_t)value_w & 15) == 0), "final_norm_fwd: value head needs a 16-byte aligned weight and an output");
  cudaStream_t st = (cudaStream_t)stream;
  int sms = sm_count();
  const int warps = FN_THREADS / 32;
  int grid = ceil_div(T, warps);
  if (grid > sms * 8) grid = sms * 8;
  const __nv_bfloat16 *xp = (const __nv_bfloat16*)x, *gp = (const __nv_bfloat16*)gamma, *wp = (const __nv_bfloat16*)value_w,
                      *bp = (const __nv_bfloat16*)value_b;
  __nv_bfloat16* yp = (__nv_bfloat16*)y;
  ProfScope ps(PROF_FINAL_NORM, st);
  if (value_w) final_norm_fwd_kernel<true><<<grid, FN_THREADS, 0, st>>>(xp, ldx, gp, eps, wp, bp, T, H, yp, ldy, rstd, value);
  else final_norm_fwd_kernel<false><<<grid, FN_THREADS, 0, st>>>(xp, ldx, gp, eps, wp, bp, T, H, yp, ldy, rstd, value);
  count_launch();
  B200_CUDA_OK(cudaGetLastError());
  return B200GRPO_OK;
}

extern "C" int b200grpo_final_norm_bwd(const void* x, int64_t ldx, const void* gamma, const float* rstd, const void* dy,
                                       int64_t lddy, const float* dvalue, const void* value_w, int T, int H, void* dx,
                                       int64_t lddx, float* dgamma, float* dvalue_w, float* dvalue_b, void* workspace,
                                       size_t workspace_bytes, void* stream) {
  if (int e = check_device_sm100()) return e;
  if (int e = fn_check(x, ldx, dx, lddx, T, H)) return e;
  B200_REQUIRE(gamma && rstd && dy && dgamma && workspace && ((uintptr_t)gamma & 15) == 0 && ((uintptr_t)dy & 15) == 0 &&
                   lddy >= H && lddy % 8 == 0, "final_norm_bwd: bad argument");
  const bool val = dvalue != nullptr;
  B200_REQUIRE(!val || (value_w && dvalue_w && ((uintptr_t)value_w & 15) == 0), "final_norm_bwd: value head needs value_w and dvalue_w");
  const int grid = fn_bwd_grid(T, H);
  const size_t part_bytes = align_up((size_t)grid * 2 * H * sizeof(float), 256);
  if (workspace_bytes < part_bytes + align_up((size_t)grid * sizeof(float), 256) || ((uintptr_t)workspace & 15)) {
    set_error("final_norm_bwd: workspace too small or misaligned");
    return B200GRPO_EWORKSPACE;
  }
  float* part = (float*)workspace;
  float* part_b = (float*)((char*)workspace + part_bytes);
  cudaStream_t st = (cudaStream_t)stream;
  const __nv_bfloat16 *xp = (const __nv_bfloat16*)x, *gp = (const __nv_bfloat16*)gamma, *dp = (const __nv_bfloat16*)dy,
                      *wp = (const __nv_bfloat16*)value_w;
  __nv_bfloat16* op = (__nv_bfloat16*)dx;
  ProfScope ps(PROF_FINAL_NORM, st);
  const size_t ring_bytes = (size_t)FN_DEPTH * 2 * H * sizeof(__nv_bfloat16);
#define FN_BWD_ONE(N, TH, V)                                                                                                    \
  do {                                                                                                                          \
    static PerDeviceOnce configured;                                                                                                   \
    if (configured.first()) {                                                                                                          \
      B200_CUDA_OK(cudaFuncSetAttribute(final_norm_bwd_kernel<N, TH, V>, cudaFuncAttributeMaxDynamicSharedMemorySize,           \
                                        FN_DEPTH * 2 * 8 * TH * N * (int)sizeof(__nv_bfloat16)));                               \
    }                                                                                                                           \
    final_norm_bwd_kernel<N, TH, V><<<grid, TH, ring_bytes, st>>>(xp, ldx, gp, rstd, dp, lddy, dvalue, wp, T, H, op, lddx, part, part_b); \
  } while (0)
#define FN_BWD(N, TH)                  \
  do {                                 \
    if (val) FN_BWD_ONE(N, TH, true);  \
    else FN_BWD_ONE(N, TH, false);     \
  } while (0)
  // every thread owns one or two 8-column chunks: 256-thread blocks (two per SM) up to H = 4096, 512-thread blocks above
  // (H in (4096, 5120]: 640 threads with ONE chunk each -- 512 threads x 2 chunks leaves three quarters of the block idle in its
  //  second round at H = 5120, the Qwen2.5-32B width)
  if (H <= 2048) FN_BWD(1, 256);
  else if (H <= 4096) FN_BWD(2, 256);
  else if (H <= 5120) FN_BWD(1, 640);
  else FN_BWD(2, 512);
#undef FN_BWD
#undef FN_BWD_ONE
  final_norm_reduce_kernel<<<ceil_div(H, FN_THREADS), FN_THREADS, 0, st>>>(part, part_b, grid, H, dgamma, val ? dvalue_w : nullptr,
                                                                           val ? dvalue_b : nullptr);
  count_launch(2);
  B200_CUDA_OK(cudaGetLastError());
  return B200GRPO_OK;
}
