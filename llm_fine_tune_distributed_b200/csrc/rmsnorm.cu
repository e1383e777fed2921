// Final RMSNorm as the producer of the fused head's X operand (first half).
//
// The reference normalises all B*T hidden rows in six ATen passes (to fp32, pow, mean, rsqrt, mul,
// cast, mul by weight -- SP/transformers/models/smollm3/modeling_smollm3.py:271-276) and the loss
// then ignores every row labelled -100.  Here the norm rides on the gather the compaction does
// anyway: one pass reads the live pre-norm rows and writes the normalised compact rows; the backward
// rides on the scatter.  HBM-bound: one warp per row, 16-byte accesses; rows of 1024 / 2048 elements
// stay in registers between the two passes, other widths re-read the row from L1.  Rounding points
// follow the reference: fp32 statistics, the normalised value rounded to bf16, then a bf16 multiply
// by the weight.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/b200sft.h"
#include "capi_common.h"

namespace b200 {
namespace {

constexpr int NORM_WARPS = 8;  // rows in flight per block

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
  return v;
}

__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
  const __nv_bfloat162* p = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const float2 t = __bfloat1622float2(p[k]);
    f[2 * k] = t.x;
    f[2 * k + 1] = t.y;
  }
}

__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  uint4 v;
  __nv_bfloat162* p = reinterpret_cast<__nv_bfloat162*>(&v);
#pragma unroll
  for (int k = 0; k < 4; ++k) p[k] = __floats2bfloat162_rn(f[2 * k], f[2 * k + 1]);
  return v;
}

__device__ __forceinline__ float round_bf16(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }

// xc[j,:] = gamma * bf16( h[row(j),:] * rsqrt(mean(h^2) + eps) ),  rstd[j] = that rsqrt.
// row(j) = row_index[j] (compaction list) or j; rows in [live, next multiple of 128) are zero-filled
// like gather_rows_kernel does (the dW GEMM contracts over whole 64-row blocks).
__global__ void __launch_bounds__(NORM_WARPS * 32)
rmsnorm_gather_kernel(const __nv_bfloat16* __restrict__ h, const __nv_bfloat16* __restrict__ gamma, float eps,
                      const int32_t* __restrict__ row_index, const int32_t* __restrict__ n_live, int N, int H,
                      __nv_bfloat16* __restrict__ xc, float* __restrict__ rstd) {
  const int live = n_live ? min(max(*n_live, 0), N) : N;
  const int padded = min((live + 127) / 128 * 128, N);
  const int lane = threadIdx.x & 31;
  const int vec = H / 8;
  const float inv_h = 1.f / static_cast<float>(H);
  const uint4* g4 = reinterpret_cast<const uint4*>(gamma);
  for (int j = blockIdx.x * NORM_WARPS + (threadIdx.x >> 5); j < padded; j += gridDim.x * NORM_WARPS) {
    uint4* dst = reinterpret_cast<uint4*>(xc + static_cast<size_t>(j) * H);
    if (j >= live) {
      for (int v = lane; v < vec; v += 32) dst[v] = make_uint4(0, 0, 0, 0);
      continue;
    }
    const int src_row = row_index ? row_index[j] : j;
    const uint4* src = reinterpret_cast<const uint4*>(h + static_cast<size_t>(src_row) * H);
    float ss = 0.f;
    for (int v = lane; v < vec; v += 32) {
      float x[8];
      unpack8(src[v], x);
#pragma unroll
      for (int k = 0; k < 8; ++k) ss = fmaf(x[k], x[k], ss);
    }
    const float r = rsqrtf(warp_sum(ss) * inv_h + eps);
    for (int v = lane; v < vec; v += 32) {
      float x[8], g[8], y[8];
      unpack8(src[v], x);
      unpack8(g4[v], g);
#pragma unroll
      for (int k = 0; k < 8; ++k) y[k] = g[k] * round_bf16(x[k] * r);
      dst[v] = pack8(y);
    }
    if (lane == 0) rstd[j] = r;
  }
}

// Same, for H = 256 * VPL: the lane's VPL vectors of the row stay in registers, so the row is read once
// and all its loads are in flight together.  Same summation order as the generic kernel (same bits).
template <int VPL>
__global__ void __launch_bounds__(NORM_WARPS * 32)
rmsnorm_gather_reg_kernel(const __nv_bfloat16* __restrict__ h, const __nv_bfloat16* __restrict__ gamma, float eps,
                          const int32_t* __restrict__ row_index, const int32_t* __restrict__ n_live, int N,
                          __nv_bfloat16* __restrict__ xc, float* __restrict__ rstd) {
  constexpr int H = VPL * 256;
  const int live = n_live ? min(max(*n_live, 0), N) : N;
  const int padded = min((live + 127) / 128 * 128, N);
  const int lane = threadIdx.x & 31;
  const uint4* g4 = reinterpret_cast<const uint4*>(gamma);
  for (int j = blockIdx.x * NORM_WARPS + (threadIdx.x >> 5); j < padded; j += gridDim.x * NORM_WARPS) {
    uint4* dst = reinterpret_cast<uint4*>(xc + static_cast<size_t>(j) * H);
    if (j >= live) {
#pragma unroll
      for (int k = 0; k < VPL; ++k) dst[lane + 32 * k] = make_uint4(0, 0, 0, 0);
      continue;
    }
    const int src_row = row_index ? row_index[j] : j;
    const uint4* src = reinterpret_cast<const uint4*>(h + static_cast<size_t>(src_row) * H);
    uint4 xv[VPL];
#pragma unroll
    for (int k = 0; k < VPL; ++k) xv[k] = src[lane + 32 * k];
    float ss = 0.f;
#pragma unroll
    for (int k = 0; k < VPL; ++k) {
      float x[8];
      unpack8(xv[k], x);
#pragma unroll
      for (int e = 0; e < 8; ++e) ss = fmaf(x[e], x[e], ss);
    }
    const float r = rsqrtf(warp_sum(ss) * (1.f / static_cast<float>(H)) + eps);
#pragma unroll
    for (int k = 0; k < VPL; ++k) {
      float x[8], g[8], y[8];
      unpack8(xv[k], x);
      unpack8(g4[lane + 32 * k], g);
#pragma unroll
      for (int e = 0; e < 8; ++e) y[e] = g[e] * round_bf16(x[e] * r);
      dst[lane + 32 * k] = pack8(y);
    }
    if (lane == 0) rstd[j] = r;
  }
}

template <int VPL>
__global__ void __launch_bounds__(NORM_WARPS * 32)
rmsnorm_scatter_bwd_reg_kernel(const __nv_bfloat16* __restrict__ dyc, const __nv_bfloat16* __restrict__ h,
                               const __nv_bfloat16* __restrict__ gamma, const float* __restrict__ rstd,
                               const int32_t* __restrict__ pos, int N, __nv_bfloat16* __restrict__ dh) {
  constexpr int H = VPL * 256;
  const int lane = threadIdx.x & 31;
  const uint4* g4 = reinterpret_cast<const uint4*>(gamma);
  for (int row = blockIdx.x * NORM_WARPS + (threadIdx.x >> 5); row < N; row += gridDim.x * NORM_WARPS) {
    uint4* dst = reinterpret_cast<uint4*>(dh + static_cast<size_t>(row) * H);
    const int j = pos ? pos[row] : row;
    if (j < 0) {
#pragma unroll
      for (int k = 0; k < VPL; ++k) dst[lane + 32 * k] = make_uint4(0, 0, 0, 0);
      continue;
    }
    const uint4* dy4 = reinterpret_cast<const uint4*>(dyc + static_cast<size_t>(j) * H);
    const uint4* h4 = reinterpret_cast<const uint4*>(h + static_cast<size_t>(row) * H);
    uint4 dv[VPL], xv[VPL];
#pragma unroll
    for (int k = 0; k < VPL; ++k) {
      dv[k] = dy4[lane + 32 * k];
      xv[k] = h4[lane + 32 * k];
    }
    const float r = rstd[j];
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < VPL; ++k) {
      float dy[8], x[8], g[8];
      unpack8(dv[k], dy);
      unpack8(xv[k], x);
      unpack8(g4[lane + 32 * k], g);
#pragma unroll
      for (int e = 0; e < 8; ++e) acc = fmaf(round_bf16(dy[e] * g[e]), x[e] * r, acc);
    }
    const float c = warp_sum(acc) * (1.f / static_cast<float>(H));
#pragma unroll
    for (int k = 0; k < VPL; ++k) {
      float dy[8], x[8], g[8], o[8];
      unpack8(dv[k], dy);
      unpack8(xv[k], x);
      unpack8(g4[lane + 32 * k], g);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float n = x[e] * r;
        o[e] = r * (round_bf16(dy[e] * g[e]) - n * c);
      }
      dst[lane + 32 * k] = pack8(o);
    }
  }
}

// dh[r,:] = rstd * (dn - n * mean(dn * n)) with n = h[r,:] * rstd, dn = bf16(dy * gamma), for the rows
// that were live (pos[r] = their compact row, or r when pos is null); zero for ignored rows.  Every
// row of dh is written exactly once.
__global__ void __launch_bounds__(NORM_WARPS * 32)
rmsnorm_scatter_bwd_kernel(const __nv_bfloat16* __restrict__ dyc, const __nv_bfloat16* __restrict__ h,
                           const __nv_bfloat16* __restrict__ gamma, const float* __restrict__ rstd,
                           const int32_t* __restrict__ pos, int N, int H, __nv_bfloat16* __restrict__ dh) {
  const int lane = threadIdx.x & 31;
  const int vec = H / 8;
  const float inv_h = 1.f / static_cast<float>(H);
  const uint4* g4 = reinterpret_cast<const uint4*>(gamma);
  for (int row = blockIdx.x * NORM_WARPS + (threadIdx.x >> 5); row < N; row += gridDim.x * NORM_WARPS) {
    uint4* dst = reinterpret_cast<uint4*>(dh + static_cast<size_t>(row) * H);
    const int j = pos ? pos[row] : row;
    if (j < 0) {
      for (int v = lane; v < vec; v += 32) dst[v] = make_uint4(0, 0, 0, 0);
      continue;
    }
    const uint4* dy4 = reinterpret_cast<const uint4*>(dyc + static_cast<size_t>(j) * H);
    const uint4* h4 = reinterpret_cast<const uint4*>(h + static_cast<size_t>(row) * H);
    const float r = rstd[j];
    float acc = 0.f;
    for (int v = lane; v < vec; v += 32) {
      float dy[8], x[8], g[8];
      unpack8(dy4[v], dy);
      unpack8(h4[v], x);
      unpack8(g4[v], g);
#pragma unroll
      for (int k = 0; k < 8; ++k) acc = fmaf(round_bf16(dy[k] * g[k]), x[k] * r, acc);
    }
    const float c = warp_sum(acc) * inv_h;
    for (int v = lane; v < vec; v += 32) {
      float dy[8], x[8], g[8], o[8];
      unpack8(dy4[v], dy);
      unpack8(h4[v], x);
      unpack8(g4[v], g);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const float n = x[k] * r;
        o[k] = r * (round_bf16(dy[k] * g[k]) - n * c);
      }
      dst[v] = pack8(o);
    }
  }
}

// dgamma partials: block (bx, by) owns columns [256*bx, 256*bx + 256) and the compact rows
// by, by + gridDim.y, ... ; its 8 warps take those rows in turn, each lane 8 consecutive columns.
// partial[by][col] = sum over the block's rows of bf16(dy * bf16(h * rstd)), warps folded in a fixed
// order (deterministic).
constexpr int DGAMMA_SPLITS = 16;

__global__ void __launch_bounds__(NORM_WARPS * 32)
rmsnorm_dgamma_partial_kernel(const __nv_bfloat16* __restrict__ dyc, const __nv_bfloat16* __restrict__ h,
                              const float* __restrict__ rstd, const int32_t* __restrict__ row_index,
                              const int32_t* __restrict__ n_live, int N, int H, float* __restrict__ partial) {
  __shared__ float s_acc[NORM_WARPS][256];
  const int live = n_live ? min(max(*n_live, 0), N) : N;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int col0 = blockIdx.x * 256 + lane * 8;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (col0 < H) {
    for (int j = blockIdx.y * NORM_WARPS + warp; j < live; j += gridDim.y * NORM_WARPS) {
      const int row = row_index ? row_index[j] : j;
      const float r = rstd[j];
      float dy[8], x[8];
      unpack8(*reinterpret_cast<const uint4*>(dyc + static_cast<size_t>(j) * H + col0), dy);
      unpack8(*reinterpret_cast<const uint4*>(h + static_cast<size_t>(row) * H + col0), x);
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] += round_bf16(dy[k] * round_bf16(x[k] * r));
    }
  }
#pragma unroll
  for (int k = 0; k < 8; ++k) s_acc[warp][lane * 8 + k] = acc[k];
  __syncthreads();
  const int col = blockIdx.x * 256 + threadIdx.x;
  if (threadIdx.x < 256 && col < H) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < NORM_WARPS; ++w) t += s_acc[w][threadIdx.x];
    partial[static_cast<size_t>(blockIdx.y) * H + col] = t;
  }
}

__global__ void rmsnorm_dgamma_finish_kernel(const float* __restrict__ partial, int H, int splits,
                                             __nv_bfloat16* __restrict__ dgamma) {
  const int col = blockIdx.x * blockDim.x + threadIdx.x;
  if (col >= H) return;
  float t = 0.f;
  for (int s = 0; s < splits; ++s) t += partial[static_cast<size_t>(s) * H + col];
  dgamma[col] = __float2bfloat16_rn(t);
}

int row_grid(int rows) {
  int dev = 0, sms = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (sms <= 0) sms = 148;
  int grid = (rows + NORM_WARPS - 1) / NORM_WARPS;
  if (grid > sms * 8) grid = sms * 8;
  return grid < 1 ? 1 : grid;
}

bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace
}  // namespace b200

using namespace b200;

extern "C" {

int b200sft_rmsnorm_gather(const void* h, const void* gamma, float eps, const int32_t* row_index,
                           const int32_t* n_live, int N, int H, void* xc, float* rstd, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (N <= 0 || H <= 0 || H % 8 != 0) return fail(B200SFT_ERR_INVALID, "rmsnorm_gather: bad shape N=%d H=%d", N, H);
  if (!h || !gamma || !xc || !rstd) return fail(B200SFT_ERR_INVALID, "null tensor pointer");
  if ((row_index == nullptr) != (n_live == nullptr))
    return fail(B200SFT_ERR_INVALID, "row_index and n_live must be given together");
  if (!aligned16(h) || !aligned16(gamma) || !aligned16(xc))
    return fail(B200SFT_ERR_INVALID, "rmsnorm_gather needs 16-byte aligned tensors");
  const auto* hp = static_cast<const __nv_bfloat16*>(h);
  const auto* gp = static_cast<const __nv_bfloat16*>(gamma);
  auto* xp = static_cast<__nv_bfloat16*>(xc);
  const int grid = row_grid(N), threads = NORM_WARPS * 32;
  switch (H) {  // rows of 256 * VPL elements stay in registers
    case 1024: rmsnorm_gather_reg_kernel<4><<<grid, threads, 0, stream>>>(hp, gp, eps, row_index, n_live, N, xp, rstd); break;
    case 2048: rmsnorm_gather_reg_kernel<8><<<grid, threads, 0, stream>>>(hp, gp, eps, row_index, n_live, N, xp, rstd); break;
    default: rmsnorm_gather_kernel<<<grid, threads, 0, stream>>>(hp, gp, eps, row_index, n_live, N, H, xp, rstd);
  }
  B200_CUDA_CHECK(cudaGetLastError());
  return B200SFT_OK;
}

size_t b200sft_rmsnorm_bwd_workspace_bytes(int H) {
  return H > 0 ? static_cast<size_t>(DGAMMA_SPLITS) * H * sizeof(float) : 0;
}

int b200sft_rmsnorm_scatter_bwd(const void* dyc, const void* h, const void* gamma, const float* rstd,
                                const int32_t* pos, const int32_t* row_index, const int32_t* n_live, int N, int H,
                                void* dh, void* dgamma, void* workspace, size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (N <= 0 || H <= 0 || H % 8 != 0)
    return fail(B200SFT_ERR_INVALID, "rmsnorm_scatter_bwd: bad shape N=%d H=%d", N, H);
  if (!dyc || !h || !gamma || !rstd) return fail(B200SFT_ERR_INVALID, "null tensor pointer");
  if (!dh && !dgamma) return fail(B200SFT_ERR_INVALID, "nothing to compute: dh and dgamma are both null");
  const bool compacted = pos || row_index || n_live;
  if (compacted && !(pos && row_index && n_live))
    return fail(B200SFT_ERR_INVALID, "pos, row_index and n_live must be given together");
  if (!aligned16(dyc) || !aligned16(h) || !aligned16(gamma) || (dh && !aligned16(dh)))
    return fail(B200SFT_ERR_INVALID, "rmsnorm_scatter_bwd needs 16-byte aligned tensors");
  if (dh) {
    const auto* dp = static_cast<const __nv_bfloat16*>(dyc);
    const auto* hp = static_cast<const __nv_bfloat16*>(h);
    const auto* gp = static_cast<const __nv_bfloat16*>(gamma);
    auto* op = static_cast<__nv_bfloat16*>(dh);
    const int grid = row_grid(N), threads = NORM_WARPS * 32;
    switch (H) {
      case 1024: rmsnorm_scatter_bwd_reg_kernel<4><<<grid, threads, 0, stream>>>(dp, hp, gp, rstd, pos, N, op); break;
      case 2048: rmsnorm_scatter_bwd_reg_kernel<8><<<grid, threads, 0, stream>>>(dp, hp, gp, rstd, pos, N, op); break;
      default: rmsnorm_scatter_bwd_kernel<<<grid, threads, 0, stream>>>(dp, hp, gp, rstd, pos, N, H, op);
    }
    B200_CUDA_CHECK(cudaGetLastError());
  }
  if (dgamma) {
    const size_t need = b200sft_rmsnorm_bwd_workspace_bytes(H);
    if (!workspace || workspace_bytes < need)
      return fail(B200SFT_ERR_WORKSPACE, "rmsnorm dgamma workspace too small: have %zu, need %zu", workspace_bytes, need);
    float* partial = static_cast<float*>(workspace);
    const dim3 grid((H + 255) / 256, DGAMMA_SPLITS);
    rmsnorm_dgamma_partial_kernel<<<grid, NORM_WARPS * 32, 0, stream>>>(
        static_cast<const __nv_bfloat16*>(dyc), static_cast<const __nv_bfloat16*>(h), rstd, row_index, n_live, N, H,
        partial);
    B200_CUDA_CHECK(cudaGetLastError());
    rmsnorm_dgamma_finish_kernel<<<(H + 255) / 256, 256, 0, stream>>>(partial, H, DGAMMA_SPLITS,
                                                                      static_cast<__nv_bfloat16*>(dgamma));
    B200_CUDA_CHECK(cudaGetLastError());
  }
  return B200SFT_OK;
}

}  // extern "C"
