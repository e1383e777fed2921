// Gradient-bucket allreduce over NVLink 5 / NVSwitch with the 1/world scale fused in.
//
// The bucket lives in symmetric memory (every rank maps every peer's copy; with NVLS also one
// multicast address that fans a store out to all copies and reduces a load across them inside the
// switch).  Each rank owns the slice [rank * n / world, (rank+1) * n / world) of the bucket:
//
//   multimem kernel : v = multimem.ld_reduce.add (fp32 accumulate) over the multicast address  -- the
//                     NVSwitch returns the sum of all ranks' copies;  v *= scale;  multimem.st v  --
//                     the switch writes the result into every rank's copy.
//   p2p kernel      : the same slice-owner scheme with explicit peer loads / stores, for boxes
//                     without multicast support.
//
// One pass over the bucket: the scale happens in registers between the reduce and the broadcast
// (single bf16 rounding of the fp32 sum * scale), instead of a separate HBM pass before the
// collective.  Neither kernel waits on another rank: the cross-rank ordering (copies complete
// before anyone reduces; results complete before anyone reads) is provided by the caller with
// symmetric-memory barriers before and after the launch, so a rank that is late can delay its peers
// but a kernel can never spin forever.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/b200sft.h"
#include "capi_common.h"

namespace b200 {

__device__ __forceinline__ float2 bf16x2_to_f2(uint32_t v) {
  return __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&v));
}
__device__ __forceinline__ uint32_t f2_to_bf16x2(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ uint32_t scale_packed(uint32_t v, float s) {
  const float2 f = bf16x2_to_f2(v);
  return f2_to_bf16x2(f.x * s, f.y * s);
}

constexpr int AR_UNROLL = 4;  // independent 16-byte multimem / peer accesses in flight per thread

__device__ __forceinline__ uint4 multimem_ld_reduce_bf16x8(const uint4* p) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st_b128(uint4* p, uint4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z),
               "r"(v.w)
               : "memory");
}

// nvec = number of 16-byte vectors (8 bf16) in the bucket; must be a multiple of `world`.
// Block-contiguous chunks (as the streaming kernels in elementwise.cu): per pass a block owns
// AR_UNROLL * blockDim consecutive vectors of this rank's slice.
__global__ void __launch_bounds__(512)
    allreduce_scale_multimem_bf16_kernel(uint4* mc, size_t nvec, int rank, int world, float scale) {
  const size_t per = nvec / world;
  uint4* base = mc + static_cast<size_t>(rank) * per;
  const size_t chunk = static_cast<size_t>(blockDim.x) * AR_UNROLL;
  for (size_t c0 = blockIdx.x * chunk; c0 < per; c0 += static_cast<size_t>(gridDim.x) * chunk) {
    uint4 v[AR_UNROLL];
#pragma unroll
    for (int u = 0; u < AR_UNROLL; ++u) {
      const size_t i = c0 + threadIdx.x + u * blockDim.x;
      if (i < per) v[u] = multimem_ld_reduce_bf16x8(base + i);
    }
#pragma unroll
    for (int u = 0; u < AR_UNROLL; ++u) {
      const size_t i = c0 + threadIdx.x + u * blockDim.x;
      if (i < per) {
        if (scale != 1.0f) {
          v[u].x = scale_packed(v[u].x, scale);
          v[u].y = scale_packed(v[u].y, scale);
          v[u].z = scale_packed(v[u].z, scale);
          v[u].w = scale_packed(v[u].w, scale);
        }
        multimem_st_b128(base + i, v[u]);
      }
    }
  }
}

constexpr int MAX_PEERS = 16;
struct PeerPtrs {
  uint4* p[MAX_PEERS];
};

__global__ void __launch_bounds__(512)
    allreduce_scale_p2p_bf16_kernel(PeerPtrs peers, size_t nvec, int rank, int world, float scale) {
  const size_t per = nvec / world;
  const size_t first = static_cast<size_t>(rank) * per;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < per;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {  // consecutive blocks take consecutive 8 KiB runs
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    uint4 v[MAX_PEERS];
#pragma unroll
    for (int r = 0; r < MAX_PEERS; ++r)  // all peer loads in flight before the first add
      if (r < world) v[r] = peers.p[r][first + i];
#pragma unroll
    for (int r = 0; r < MAX_PEERS; ++r) {
      if (r < world) {
        const float2 f0 = bf16x2_to_f2(v[r].x), f1 = bf16x2_to_f2(v[r].y), f2 = bf16x2_to_f2(v[r].z),
                     f3 = bf16x2_to_f2(v[r].w);
        acc[0] += f0.x; acc[1] += f0.y; acc[2] += f1.x; acc[3] += f1.y;
        acc[4] += f2.x; acc[5] += f2.y; acc[6] += f3.x; acc[7] += f3.y;
      }
    }
    uint4 o;
    o.x = f2_to_bf16x2(acc[0] * scale, acc[1] * scale);
    o.y = f2_to_bf16x2(acc[2] * scale, acc[3] * scale);
    o.z = f2_to_bf16x2(acc[4] * scale, acc[5] * scale);
    o.w = f2_to_bf16x2(acc[6] * scale, acc[7] * scale);
#pragma unroll
    for (int r = 0; r < MAX_PEERS; ++r)
      if (r < world) peers.p[r][first + i] = o;
  }
}

}  // namespace b200

using namespace b200;

namespace {
int ar_grid(size_t per_rank_vecs, int threads) {
  int dev = 0, sms = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (sms <= 0) sms = 148;
  size_t blocks = (per_rank_vecs + threads - 1) / threads;
  const size_t cap = static_cast<size_t>(sms) * 2;  // 1024 threads/SM: leaves room for the backward kernels it overlaps with
  if (blocks > cap) blocks = cap;
  if (blocks == 0) blocks = 1;
  return static_cast<int>(blocks);
}
}  // namespace

extern "C" {

int b200sft_allreduce_scale_bf16_multimem(void* multicast_ptr, size_t n, int rank, int world, float scale,
                                          void* stream) {
  if (!multicast_ptr) return fail(B200SFT_ERR_INVALID, "null multicast pointer (no NVLS support?)");
  if (world <= 0 || rank < 0 || rank >= world) return fail(B200SFT_ERR_INVALID, "bad rank %d / world %d", rank, world);
  if (n % (8 * static_cast<size_t>(world)) != 0 || (reinterpret_cast<uintptr_t>(multicast_ptr) & 15) != 0)
    return fail(B200SFT_ERR_INVALID, "bucket must be 16-byte aligned with a length that is a multiple of 8*world (n=%zu)", n);
  if (n == 0) return B200SFT_OK;
  const size_t nvec = n / 8;
  allreduce_scale_multimem_bf16_kernel<<<ar_grid(nvec / world, 512), 512, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<uint4*>(multicast_ptr), nvec, rank, world, scale);
  B200_CUDA_CHECK(cudaGetLastError());
  return B200SFT_OK;
}

int b200sft_allreduce_scale_bf16_p2p(const uint64_t* peer_ptrs_host, size_t n, int rank, int world, float scale,
                                     void* stream) {
  if (!peer_ptrs_host) return fail(B200SFT_ERR_INVALID, "null peer pointer table");
  if (world <= 0 || world > MAX_PEERS || rank < 0 || rank >= world)
    return fail(B200SFT_ERR_INVALID, "bad rank %d / world %d (at most %d peers)", rank, world, MAX_PEERS);
  if (n % (8 * static_cast<size_t>(world)) != 0)
    return fail(B200SFT_ERR_INVALID, "bucket length must be a multiple of 8*world (n=%zu)", n);
  if (n == 0) return B200SFT_OK;
  PeerPtrs pp = {};
  for (int r = 0; r < world; ++r) {
    if (peer_ptrs_host[r] == 0 || (peer_ptrs_host[r] & 15) != 0) return fail(B200SFT_ERR_INVALID, "bad peer pointer %d", r);
    pp.p[r] = reinterpret_cast<uint4*>(peer_ptrs_host[r]);
  }
  const size_t nvec = n / 8;
  allreduce_scale_p2p_bf16_kernel<<<ar_grid(nvec / world, 512), 512, 0, static_cast<cudaStream_t>(stream)>>>(
      pp, nvec, rank, world, scale);
  B200_CUDA_CHECK(cudaGetLastError());
  return B200SFT_OK;
}

}  // extern "C"
