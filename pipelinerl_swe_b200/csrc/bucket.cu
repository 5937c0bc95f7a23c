// K6: gather many parameters into one contiguous bf16 bucket (and scatter back), so the
// trainer->actor weight update of /root/reference/pipelinerl/finetune_loop.py:236-248 and
// vllm1.py:81-94 becomes a few large NCCL broadcasts instead of one launch + one
// allocation per parameter.  HBM-bound byte copy: 4 B/element (bf16 in + bf16 out), or
// 6 B/element when the source holds fp32 master weights (cast fused into the pack).
// Every tensor starts on a 16-byte boundary inside the bucket (offsets rounded up to 8
// elements) so both sides move 128-bit vectors.
//
// Multicast pack (b200grpo_bucket_pack_multicast): the bucket address is an NVSwitch multicast mapping of a
// symmetric buffer that every rank of the update group owns, and the pack stores with multimem.st -- ONE kernel
// reads the parameters once and the switch delivers the bf16 bytes into every receiver's copy: pack and
// broadcast are the same pass, the sender's NVLink egress carries the payload once whatever the group size.
//
// Scatter + relay (b200grpo_bucket_pack_scatter / b200grpo_bucket_relay_multicast): the design that lets a 1 -> N broadcast
// run at the line rate of ONE copy whatever N is.  The bucket is cut into stripes; the sender's pack kernel stores stripe j
// straight into receiver (j mod N)'s copy of the symmetric slot over P2P NVLink stores (egress: the payload once, spread
// over N peers -- no staging buffer on the sender at all), then every receiver re-emits the stripes it owns with
// multimem.st, the switch replicating them into all other copies (egress per receiver: payload / N; ingress: the payload
// once).  Sender egress and receiver ingress are both exactly one payload, and the two phases of consecutive transfers
// overlap (two slots).
#include "common.cuh"
#include <vector>

namespace b200grpo {

constexpr int BK_THREADS = 256;
constexpr int BK_CHUNK = 8192;  // elements per work item (32 per thread)

struct BucketEntry {
  void* ptr;            // parameter storage (src for pack, dst for unpack)
  int64_t numel;
  int64_t bucket_off;   // element offset inside the bucket (multiple of 8)
  int64_t chunk_start;  // first work item of this tensor
  int32_t is_f32;
  int32_t pad;
};

__device__ __forceinline__ int find_entry(const BucketEntry* tab, int n, int64_t chunk) {
  int lo = 0, hi = n - 1;
  while (lo < hi) {  // last entry with chunk_start <= chunk
    const int mid = (lo + hi + 1) >> 1;
    if (tab[mid].chunk_start <= chunk) lo = mid; else hi = mid - 1;
  }
  return lo;
}

// 16-byte store to an NVSwitch multicast address: lands in the buffer of every rank bound to the mapping
__device__ __forceinline__ void multimem_st_16(void* mc_addr, uint4 v) {
  asm volatile("multimem.st.weak.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc_addr), "f"(__uint_as_float(v.x)),
               "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)), "f"(__uint_as_float(v.w)) : "memory");
}

// multicast pack: every store is a whole 16-byte vector (tails are zero-padded: bucket offsets are multiples of 8)
__global__ void __launch_bounds__(BK_THREADS) bucket_multicast_kernel(const BucketEntry* __restrict__ tab, int n,
                                                                      int64_t total_chunks, __nv_bfloat16* mc_bucket) {
  for (int64_t chunk = blockIdx.x; chunk < total_chunks; chunk += gridDim.x) {
    const BucketEntry e = tab[find_entry(tab, n, chunk)];
    const int64_t begin = (chunk - e.chunk_start) * BK_CHUNK;
    const int64_t end = min(begin + (int64_t)BK_CHUNK, e.numel);
    __nv_bfloat16* b = mc_bucket + e.bucket_off;
    const bool vec_ok = ((uintptr_t)e.ptr & 15) == 0;
    for (int64_t i = begin + threadIdx.x * 8LL; i < end; i += BK_THREADS * 8LL) {
      uint4 out;
      if (i + 8 <= end && vec_ok) {
        if (!e.is_f32) {
          out = __ldg((const uint4*)((const __nv_bfloat16*)e.ptr + i));
        } else {
          const float* p = (const float*)e.ptr;
          const float4 x = __ldg((const float4*)(p + i)), y = __ldg((const float4*)(p + i + 4));
          __nv_bfloat162 o[4] = {__floats2bfloat162_rn(x.x, x.y), __floats2bfloat162_rn(x.z, x.w),
                                 __floats2bfloat162_rn(y.x, y.y), __floats2bfloat162_rn(y.z, y.w)};
          out = *(uint4*)o;
        }
      } else {  // unaligned source or the last partial vector of the tensor
        __nv_bfloat16 o[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const int64_t j = i + k;
          o[k] = j < end ? (e.is_f32 ? __float2bfloat16_rn(((const float*)e.ptr)[j]) : ((const __nv_bfloat16*)e.ptr)[j])
                         : __float2bfloat16_rn(0.f);
        }
        out = *(uint4*)o;
      }
      multimem_st_16(b + i, out);
    }
  }
  __threadfence_system();  // the stores are performed at every peer before the grid is reported complete
}

// Scatter pack: like the multicast pack, but each 16-byte vector goes to exactly ONE peer -- the owner of its stripe --
// with a plain st.global through the peer's P2P mapping.  peers[q] = base of peer q's copy of the slot.
__global__ void __launch_bounds__(BK_THREADS) bucket_scatter_kernel(const BucketEntry* __restrict__ tab, int n, int64_t total_chunks,
                                                                    __nv_bfloat16* const* __restrict__ peers, int n_peers,
                                                                    int64_t stripe_elems) {
  for (int64_t chunk = blockIdx.x; chunk < total_chunks; chunk += gridDim.x) {
    const BucketEntry e = tab[find_entry(tab, n, chunk)];
    const int64_t begin = (chunk - e.chunk_start) * BK_CHUNK;
    const int64_t end = min(begin + (int64_t)BK_CHUNK, e.numel);
    const bool vec_ok = ((uintptr_t)e.ptr & 15) == 0;
    for (int64_t i = begin + threadIdx.x * 8LL; i < end; i += BK_THREADS * 8LL) {
      uint4 out;
      if (i + 8 <= end && vec_ok) {
        if (!e.is_f32) {
          out = __ldg((const uint4*)((const __nv_bfloat16*)e.ptr + i));
        } else {
          const float* p = (const float*)e.ptr;
          const float4 x = __ldg((const float4*)(p + i)), y = __ldg((const float4*)(p + i + 4));
          __nv_bfloat162 o[4] = {__floats2bfloat162_rn(x.x, x.y), __floats2bfloat162_rn(x.z, x.w),
                                 __floats2bfloat162_rn(y.x, y.y), __floats2bfloat162_rn(y.z, y.w)};
          out = *(uint4*)o;
        }
      } else {  // unaligned source or the last partial vector of the tensor (zero-padded to the 16-byte boundary)
        __nv_bfloat16 o[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const int64_t j = i + k;
          o[k] = j < end ? (e.is_f32 ? __float2bfloat16_rn(((const float*)e.ptr)[j]) : ((const __nv_bfloat16*)e.ptr)[j])
                         : __float2bfloat16_rn(0.f);
        }
        out = *(uint4*)o;
      }
      const int64_t off = e.bucket_off + i;                       // multiple of 8; stripes are multiples of 8 elements
      __nv_bfloat16* dst = peers[(off / stripe_elems) % n_peers];
      *(uint4*)(dst + off) = out;
    }
  }
  __threadfence_system();  // the stores are performed at every peer before the grid is reported complete
}

// Relay: this receiver re-emits the stripes it owns (stripe j with j % n_peers == peer_index) from its local copy of the
// slot to the multicast mapping of the same slot.  16 B per thread per trip, four loads in flight.
constexpr int RL_UNROLL = 4;
__global__ void __launch_bounds__(BK_THREADS) bucket_relay_kernel(const uint4* local, uint4* mc, int64_t n_vec,
                                                                  int peer_index, int n_peers, int64_t stripe_vec) {
  // enumerate only the owned vectors: owned index o -> stripe (o / stripe_vec) * n_peers + peer_index
  const int64_t n_stripes = (n_vec + stripe_vec - 1) / stripe_vec;
  const int64_t owned_stripes = n_stripes > peer_index ? (n_stripes - peer_index + n_peers - 1) / n_peers : 0;
  const int64_t n_owned = owned_stripes * stripe_vec;
  const int64_t stride = (int64_t)gridDim.x * BK_THREADS;
  for (int64_t o = (int64_t)blockIdx.x * BK_THREADS + threadIdx.x; o < n_owned; o += stride * RL_UNROLL) {
    uint4 v[RL_UNROLL];
    int64_t idx[RL_UNROLL];
#pragma unroll
    for (int u = 0; u < RL_UNROLL; ++u) {
      const int64_t oo = o + u * stride;
      const int64_t st = oo / stripe_vec;
      idx[u] = (st * n_peers + peer_index) * stripe_vec + (oo - st * stripe_vec);
      if (oo >= n_owned || idx[u] >= n_vec) idx[u] = -1;
      else v[u] = local[idx[u]];
    }
#pragma unroll
    for (int u = 0; u < RL_UNROLL; ++u)
      if (idx[u] >= 0) multimem_st_16(mc + idx[u], v[u]);
  }
  __threadfence_system();
}

// Relay over plain P2P stores instead of the multicast mapping: the owned stripes go to every OTHER receiver's copy with
// st.global (egress per receiver: payload * (N-1)/N).  Same enumeration as bucket_relay_kernel.
constexpr int RL_MAX_PEERS = 16;
struct RelayPeers { uint4* p[RL_MAX_PEERS]; };
__global__ void __launch_bounds__(BK_THREADS) bucket_relay_p2p_kernel(const uint4* local, const RelayPeers peers, int64_t n_vec,
                                                                      int peer_index, int n_peers, int64_t stripe_vec) {
  const int64_t n_stripes = (n_vec + stripe_vec - 1) / stripe_vec;
  const int64_t owned_stripes = n_stripes > peer_index ? (n_stripes - peer_index + n_peers - 1) / n_peers : 0;
  const int64_t n_owned = owned_stripes * stripe_vec;
  const int64_t stride = (int64_t)gridDim.x * BK_THREADS;
  for (int64_t o = (int64_t)blockIdx.x * BK_THREADS + threadIdx.x; o < n_owned; o += stride * RL_UNROLL) {
    uint4 v[RL_UNROLL];
    int64_t idx[RL_UNROLL];
#pragma unroll
    for (int u = 0; u < RL_UNROLL; ++u) {
      const int64_t oo = o + u * stride;
      const int64_t st = oo / stripe_vec;
      idx[u] = (st * n_peers + peer_index) * stripe_vec + (oo - st * stripe_vec);
      if (oo >= n_owned || idx[u] >= n_vec) idx[u] = -1;
      else v[u] = local[idx[u]];
    }
    for (int q = 0; q < n_peers; ++q) {
      if (q == peer_index) continue;
#pragma unroll
      for (int u = 0; u < RL_UNROLL; ++u)
        if (idx[u] >= 0) peers.p[q][idx[u]] = v[u];
    }
  }
  __threadfence_system();
}

template <bool PACK>
__global__ void __launch_bounds__(BK_THREADS) bucket_kernel(const BucketEntry* __restrict__ tab, int n,
                                                            int64_t total_chunks, __nv_bfloat16* bucket) {
  for (int64_t chunk = blockIdx.x; chunk < total_chunks; chunk += gridDim.x) {
    const BucketEntry e = tab[find_entry(tab, n, chunk)];
    const int64_t begin = (chunk - e.chunk_start) * BK_CHUNK;
    const int64_t end = min(begin + (int64_t)BK_CHUNK, e.numel);
    __nv_bfloat16* b = bucket + e.bucket_off;
    const bool vec_ok = ((uintptr_t)e.ptr & 15) == 0;
    if (!e.is_f32) {
      __nv_bfloat16* p = (__nv_bfloat16*)e.ptr;
      const int64_t vend = vec_ok ? begin + ((end - begin) & ~7LL) : begin;
      for (int64_t i = begin + threadIdx.x * 8LL; i < vend; i += BK_THREADS * 8LL) {
        if (PACK) *(uint4*)(b + i) = __ldg((const uint4*)(p + i));
        else *(uint4*)(p + i) = __ldg((const uint4*)(b + i));
      }
      for (int64_t i = vend + threadIdx.x; i < end; i += BK_THREADS) {
        if (PACK) b[i] = p[i]; else p[i] = b[i];
      }
    } else {
      float* p = (float*)e.ptr;  // fp32 master weights: cast while packing; unpack widens
      const int64_t vend = vec_ok ? begin + ((end - begin) & ~7LL) : begin;
      for (int64_t i = begin + threadIdx.x * 8LL; i < vend; i += BK_THREADS * 8LL) {
        if (PACK) {
          const float4 x = __ldg((const float4*)(p + i)), y = __ldg((const float4*)(p + i + 4));
          __nv_bfloat162 o[4] = {__floats2bfloat162_rn(x.x, x.y), __floats2bfloat162_rn(x.z, x.w),
                                 __floats2bfloat162_rn(y.x, y.y), __floats2bfloat162_rn(y.z, y.w)};
          *(uint4*)(b + i) = *(uint4*)o;
        } else {
          const uint4 raw = __ldg((const uint4*)(b + i));
          const __nv_bfloat162* o = (const __nv_bfloat162*)&raw;
          const float2 f0 = __bfloat1622float2(o[0]), f1 = __bfloat1622float2(o[1]);
          const float2 f2 = __bfloat1622float2(o[2]), f3 = __bfloat1622float2(o[3]);
          *(float4*)(p + i) = make_float4(f0.x, f0.y, f1.x, f1.y);
          *(float4*)(p + i + 4) = make_float4(f2.x, f2.y, f3.x, f3.y);
        }
      }
      for (int64_t i = vend + threadIdx.x; i < end; i += BK_THREADS) {
        if (PACK) b[i] = __float2bfloat16_rn(p[i]); else p[i] = __bfloat162float(b[i]);
      }
    }
  }
}

static int run_bucket(int mode, void* const* ptrs, const int64_t* numel, const uint8_t* is_f32, int n, void* bucket,
                      int64_t capacity, void* workspace, size_t workspace_bytes, void* stream, void* const* peers = nullptr,
                      int n_peers = 0, int64_t stripe_elems = 0) {
  if (int e = check_device_sm100()) return e;
  B200_REQUIRE(n >= 0, "bucket: n_tensors < 0");
  if (n == 0) return B200GRPO_OK;
  B200_REQUIRE(ptrs && numel && (bucket || mode == 3) && workspace, "bucket: null pointer");
  B200_REQUIRE(((uintptr_t)bucket & 15) == 0, "bucket: bucket must be 16-byte aligned");
  const size_t tab_bytes = b200grpo_bucket_workspace(n);
  if (workspace_bytes < tab_bytes + (mode == 3 ? align_up(sizeof(void*) * (size_t)n_peers, 256) : 0)) {
    set_error("bucket: workspace too small");
    return B200GRPO_EWORKSPACE;
  }
  std::vector<BucketEntry> tab(n);
  int64_t off = 0, chunks = 0;
  for (int i = 0; i < n; ++i) {
    B200_REQUIRE(numel[i] >= 0 && (numel[i] == 0 || ptrs[i]), "bucket: tensor %d invalid", i);
    tab[i] = BucketEntry{ptrs[i], numel[i], off, chunks, is_f32 ? (int32_t)is_f32[i] : 0, 0};
    off += (numel[i] + 7) / 8 * 8;
    chunks += (numel[i] + BK_CHUNK - 1) / BK_CHUNK;
  }
  B200_REQUIRE(capacity < 0 || off <= capacity, "bucket: needs %lld elements, capacity %lld", (long long)off, (long long)capacity);
  if (chunks == 0) return B200GRPO_OK;
  cudaStream_t st = (cudaStream_t)stream;
  B200_CUDA_OK(cudaMemcpyAsync(workspace, tab.data(), sizeof(BucketEntry) * n, cudaMemcpyHostToDevice, st));
  __nv_bfloat16** peers_dev = (__nv_bfloat16**)((char*)workspace + tab_bytes);
  if (mode == 3) B200_CUDA_OK(cudaMemcpyAsync(peers_dev, peers, sizeof(void*) * n_peers, cudaMemcpyHostToDevice, st));
  int64_t grid = chunks;
  const int64_t cap = (int64_t)sm_count() * 8;
  if (grid > cap) grid = cap;
  ProfScope ps(PROF_BUCKET, st);
  count_launch();
  if (mode == 3) bucket_scatter_kernel<<<(int)grid, BK_THREADS, 0, st>>>((const BucketEntry*)workspace, n, chunks, peers_dev, n_peers, stripe_elems);
  else if (mode == 2) bucket_multicast_kernel<<<(int)grid, BK_THREADS, 0, st>>>((const BucketEntry*)workspace, n, chunks, (__nv_bfloat16*)bucket);
  else if (mode == 1) bucket_kernel<true><<<(int)grid, BK_THREADS, 0, st>>>((const BucketEntry*)workspace, n, chunks, (__nv_bfloat16*)bucket);
  else bucket_kernel<false><<<(int)grid, BK_THREADS, 0, st>>>((const BucketEntry*)workspace, n, chunks, (__nv_bfloat16*)bucket);
  B200_CUDA_OK(cudaGetLastError());
  return B200GRPO_OK;
}

}  // namespace b200grpo

using namespace b200grpo;

extern "C" size_t b200grpo_bucket_workspace(int n_tensors) {
  return align_up(sizeof(BucketEntry) * (size_t)(n_tensors > 0 ? n_tensors : 1), 256);
}

extern "C" int64_t b200grpo_bucket_numel(const int64_t* numel_host, int n_tensors) {
  int64_t off = 0;
  for (int i = 0; i < n_tensors; ++i) off += (numel_host[i] + 7) / 8 * 8;
  return off;
}

extern "C" int b200grpo_bucket_pack(const void* const* src_ptrs_host, const int64_t* numel_host,
                                    const uint8_t* src_is_f32_host, int n_tensors, void* bucket_bf16,
                                    int64_t bucket_capacity, void* workspace, size_t workspace_bytes, void* stream) {
  return run_bucket(1, (void* const*)src_ptrs_host, numel_host, src_is_f32_host, n_tensors, bucket_bf16,
                    bucket_capacity, workspace, workspace_bytes, stream);
}

extern "C" int b200grpo_bucket_unpack(const void* bucket_bf16, void* const* dst_ptrs_host, const int64_t* numel_host,
                                      const uint8_t* dst_is_f32_host, int n_tensors, void* workspace,
                                      size_t workspace_bytes, void* stream) {
  return run_bucket(0, dst_ptrs_host, numel_host, dst_is_f32_host, n_tensors, (void*)bucket_bf16, -1, workspace,
                    workspace_bytes, stream);
}

extern "C" int b200grpo_bucket_pack_multicast(const void* const* src_ptrs_host, const int64_t* numel_host,
                                              const uint8_t* src_is_f32_host, int n_tensors, void* bucket_multicast_bf16,
                                              int64_t bucket_capacity, void* workspace, size_t workspace_bytes, void* stream) {
  return run_bucket(2, (void* const*)src_ptrs_host, numel_host, src_is_f32_host, n_tensors, bucket_multicast_bf16,
                    bucket_capacity, workspace, workspace_bytes, stream);
}

extern "C" size_t b200grpo_bucket_scatter_workspace(int n_tensors, int n_peers) {
  return b200grpo_bucket_workspace(n_tensors) + align_up(sizeof(void*) * (size_t)(n_peers > 0 ? n_peers : 1), 256);
}

extern "C" int b200grpo_bucket_pack_scatter(const void* const* src_ptrs_host, const int64_t* numel_host,
                                            const uint8_t* src_is_f32_host, int n_tensors, void* const* peer_buckets_host,
                                            int n_peers, int64_t stripe_elems, int64_t bucket_capacity, void* workspace,
                                            size_t workspace_bytes, void* stream) {
  B200_REQUIRE(peer_buckets_host && n_peers >= 1 && n_peers <= 64, "bucket_pack_scatter: n_peers=%d", n_peers);
  B200_REQUIRE(stripe_elems >= 8 && stripe_elems % 8 == 0, "bucket_pack_scatter: stripe must be a multiple of 8 elements");
  for (int q = 0; q < n_peers; ++q)
    B200_REQUIRE(peer_buckets_host[q] && ((uintptr_t)peer_buckets_host[q] & 15) == 0, "bucket_pack_scatter: peer %d pointer", q);
  return run_bucket(3, (void* const*)src_ptrs_host, numel_host, src_is_f32_host, n_tensors, nullptr, bucket_capacity, workspace,
                    workspace_bytes, stream, peer_buckets_host, n_peers, stripe_elems);
}

extern "C" int b200grpo_bucket_relay_multicast(const void* local_bucket, void* bucket_multicast, int64_t numel, int peer_index,
                                               int n_peers, int64_t stripe_elems, int ctas, void* stream) {
  if (int e = check_device_sm100()) return e;
  B200_REQUIRE(local_bucket && bucket_multicast && (((uintptr_t)local_bucket | (uintptr_t)bucket_multicast) & 15) == 0,
               "bucket_relay: 16-byte aligned pointers");
  B200_REQUIRE(numel >= 0 && numel % 8 == 0 && stripe_elems >= 8 && stripe_elems % 8 == 0, "bucket_relay: numel / stripe %% 8");
  B200_REQUIRE(n_peers >= 1 && peer_index >= 0 && peer_index < n_peers, "bucket_relay: peer %d of %d", peer_index, n_peers);
  if (numel == 0) return B200GRPO_OK;
  const int64_t n_vec = numel / 8, stripe_vec = stripe_elems / 8;
  const int64_t own = (n_vec / n_peers) + stripe_vec;
  int64_t grid = (own + (int64_t)BK_THREADS * RL_UNROLL - 1) / ((int64_t)BK_THREADS * RL_UNROLL);
  const int64_t cap = ctas > 0 ? ctas : (int64_t)sm_count() * 4;
  if (grid > cap) grid = cap;
  if (grid < 1) grid = 1;
  cudaStream_t st = (cudaStream_t)stream;
  ProfScope ps(PROF_BUCKET, st);
  count_launch();
  bucket_relay_kernel<<<(int)grid, BK_THREADS, 0, st>>>((const uint4*)local_bucket, (uint4*)bucket_multicast, n_vec, peer_index,
                                                        n_peers, stripe_vec);
  B200_CUDA_OK(cudaGetLastError());
  return B200GRPO_OK;
}

extern "C" int b200grpo_bucket_relay_p2p(const void* local_bucket, void* const* peer_buckets_host, int64_t numel, int peer_index,
                                         int n_peers, int64_t stripe_elems, int ctas, void* stream) {
  if (int e = check_device_sm100()) return e;
  B200_REQUIRE(local_bucket && peer_buckets_host && ((uintptr_t)local_bucket & 15) == 0, "bucket_relay_p2p: pointers");
  B200_REQUIRE(numel >= 0 && numel % 8 == 0 && stripe_elems >= 8 && stripe_elems % 8 == 0, "bucket_relay_p2p: numel / stripe %% 8");
  B200_REQUIRE(n_peers >= 1 && n_peers <= RL_MAX_PEERS && peer_index >= 0 && peer_index < n_peers, "bucket_relay_p2p: peer %d of %d", peer_index, n_peers);
  if (numel == 0 || n_peers == 1) return B200GRPO_OK;
  RelayPeers peers = {};
  for (int q = 0; q < n_peers; ++q) {
    B200_REQUIRE(peer_buckets_host[q] && ((uintptr_t)peer_buckets_host[q] & 15) == 0, "bucket_relay_p2p: peer %d pointer", q);
    peers.p[q] = (uint4*)peer_buckets_host[q];
  }
  const int64_t n_vec = numel / 8, stripe_vec = stripe_elems / 8;
  const int64_t own = (n_vec / n_peers) + stripe_vec;
  int64_t grid = (own + (int64_t)BK_THREADS * RL_UNROLL - 1) / ((int64_t)BK_THREADS * RL_UNROLL);
  const int64_t cap = ctas > 0 ? ctas : (int64_t)sm_count() * 4;
  if (grid > cap) grid = cap;
  if (grid < 1) grid = 1;
  cudaStream_t st = (cudaStream_t)stream;
  ProfScope ps(PROF_BUCKET, st);
  count_launch();
  bucket_relay_p2p_kernel<<<(int)grid, BK_THREADS, 0, st>>>((const uint4*)local_bucket, peers, n_vec, peer_index, n_peers, stripe_vec);
  B200_CUDA_OK(cudaGetLastError());
  return B200GRPO_OK;
}
