// Step tail of the SFT step on bf16 parameters: global gradient-norm clip + AdamW, as two
// multi-tensor HBM-bound kernels.
//
//   reference: accelerator.clip_grad_norm_(params, 1.0)  (training.py:264 -> SP/transformers/trainer.py:2489-2499
//              -> torch.nn.utils.clip_grad_norm_) followed by AdamW.step (torch.optim.AdamW on bf16 params).
//   here:      grad_sumsq kernel (one read of the gradients, fixed-order two-level reduction) and an
//              AdamW kernel that applies the clip coefficient while it reads the gradient -- the
//              gradients are read twice and never rewritten (the reference reads them three times and
//              writes them once).  Per element: 2 B (sumsq) + 8 B read + 6 B written (AdamW).
//
// The update follows torch's fused AdamW arithmetic for bf16 tensors (fp32 math, one rounding per
// stored value):  p -= lr*wd*p;  m = b1*m + (1-b1)*g;  v = b2*v + (1-b2)*g*g;
//                 p -= (lr / (1-b1^t)) * m / (sqrt(v)/sqrt(1-b2^t) + eps)
// with g = bf16(grad * clip_coef), clip_coef = min(1, max_norm / (||grad|| + 1e-6)) -- the clipped
// gradient is rounded to bf16 because the reference clips in place before the optimizer reads it.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "../../include/b200sft.h"
#include "capi_common.h"

namespace b200 {

constexpr int MT_MAX_TENSORS = 48;          // tensors per launch (pointer table travels as a kernel argument)
constexpr int MT_CHUNK = 8192;              // elements per block: 16 KiB of bf16
constexpr int MT_THREADS = 256;

struct TensorTable {
  void* p[MT_MAX_TENSORS];
  void* g[MT_MAX_TENSORS];
  void* m[MT_MAX_TENSORS];
  void* v[MT_MAX_TENSORS];
  unsigned long long numel[MT_MAX_TENSORS];
  int first_block[MT_MAX_TENSORS + 1];  // prefix sum of ceil(numel / MT_CHUNK)
  int n;
};

__device__ __forceinline__ int find_tensor(const TensorTable& t, int block) {
  int i = 0;
  while (i + 1 < t.n && block >= t.first_block[i + 1]) ++i;
  return i;
}

// partial[block] = sum of g^2 over the block's chunk (fp32 within the thread, fp64 across the block)
__global__ void __launch_bounds__(MT_THREADS) grad_sumsq_kernel(const __grid_constant__ TensorTable t, int block_offset,
                                                                double* __restrict__ partial) {
  const int ti = find_tensor(t, blockIdx.x);
  const size_t start = static_cast<size_t>(blockIdx.x - t.first_block[ti]) * MT_CHUNK;
  const size_t n = t.numel[ti];
  const size_t end = start + MT_CHUNK < n ? start + MT_CHUNK : n;
  const __nv_bfloat16* g = static_cast<const __nv_bfloat16*>(t.g[ti]);
  float acc = 0.f;
  const bool vec = (reinterpret_cast<uintptr_t>(g) & 15) == 0;
  if (vec && end - start == MT_CHUNK) {
    const uint4* g4 = reinterpret_cast<const uint4*>(g + start);
#pragma unroll
    for (int u = 0; u < MT_CHUNK / 8 / MT_THREADS; ++u) {
      uint4 x = g4[threadIdx.x + u * MT_THREADS];
      const uint32_t w[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[k]));
        acc = fmaf(f.x, f.x, acc);
        acc = fmaf(f.y, f.y, acc);
      }
    }
  } else {
    for (size_t i = start + threadIdx.x; i < end; i += MT_THREADS) {
      const float f = __bfloat162float(g[i]);
      acc = fmaf(f, f, acc);
    }
  }
  __shared__ double s[MT_THREADS];
  s[threadIdx.x] = static_cast<double>(acc);
  __syncthreads();
  for (int off = MT_THREADS >> 1; off > 0; off >>= 1) {
    if (threadIdx.x < off) s[threadIdx.x] += s[threadIdx.x + off];
    __syncthreads();
  }
  if (threadIdx.x == 0) partial[block_offset + blockIdx.x] = s[0];
}

// One block: total_norm = sqrt(sum partial[0..n)) in fixed order.
__global__ void __launch_bounds__(1024) grad_norm_finish_kernel(const double* __restrict__ partial, int n,
                                                                float* __restrict__ total_norm) {
  __shared__ double s[1024];
  double acc = 0.0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) acc += partial[i];
  s[threadIdx.x] = acc;
  __syncthreads();
  for (int off = blockDim.x >> 1; off > 0; off >>= 1) {
    if (threadIdx.x < off) s[threadIdx.x] += s[threadIdx.x + off];
    __syncthreads();
  }
  if (threadIdx.x == 0) *total_norm = static_cast<float>(sqrt(s[0]));
}

struct AdamArgs {
  float lr, beta1, beta2, eps, weight_decay, bias1, bias2_sqrt, max_norm;
};

__device__ __forceinline__ void adamw_one(float& p, float g, float& m, float& v, const AdamArgs& a) {
  if (a.weight_decay != 0.f) p -= a.lr * a.weight_decay * p;
  m = fmaf(a.beta1, m, fmaf(-a.beta1, g, g));
  const float gg = g * g;
  v = fmaf(a.beta2, v, fmaf(-a.beta2, gg, gg));
  const float denom = sqrtf(v) / a.bias2_sqrt + a.eps;
  p -= (a.lr / a.bias1) * m / denom;
}

__global__ void __launch_bounds__(MT_THREADS)
    adamw_clip_bf16_kernel(const __grid_constant__ TensorTable t, const AdamArgs a, const float* __restrict__ total_norm) {
  float coef = 1.f;
  if (a.max_norm > 0.f && total_norm != nullptr) coef = fminf(1.f, a.max_norm / (__ldg(total_norm) + 1e-6f));
  const int ti = find_tensor(t, blockIdx.x);
  const size_t start = static_cast<size_t>(blockIdx.x - t.first_block[ti]) * MT_CHUNK;
  const size_t n = t.numel[ti];
  const size_t end = start + MT_CHUNK < n ? start + MT_CHUNK : n;
  __nv_bfloat16* P = static_cast<__nv_bfloat16*>(t.p[ti]);
  const __nv_bfloat16* G = static_cast<const __nv_bfloat16*>(t.g[ti]);
  __nv_bfloat16* M = static_cast<__nv_bfloat16*>(t.m[ti]);
  __nv_bfloat16* V = static_cast<__nv_bfloat16*>(t.v[ti]);
  const bool vec = ((reinterpret_cast<uintptr_t>(P) | reinterpret_cast<uintptr_t>(G) | reinterpret_cast<uintptr_t>(M) |
                     reinterpret_cast<uintptr_t>(V)) & 15) == 0;
  if (vec && end - start == MT_CHUNK) {
#pragma unroll 2
    for (int u = 0; u < MT_CHUNK / 8 / MT_THREADS; ++u) {
      const size_t i8 = start / 8 + threadIdx.x + u * MT_THREADS;
      uint4 p4 = reinterpret_cast<uint4*>(P)[i8];
      const uint4 g4 = reinterpret_cast<const uint4*>(G)[i8];
      uint4 m4 = reinterpret_cast<uint4*>(M)[i8];
      uint4 v4 = reinterpret_cast<uint4*>(V)[i8];
      uint32_t* pw = reinterpret_cast<uint32_t*>(&p4);
      const uint32_t* gw = reinterpret_cast<const uint32_t*>(&g4);
      uint32_t* mw = reinterpret_cast<uint32_t*>(&m4);
      uint32_t* vw = reinterpret_cast<uint32_t*>(&v4);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        float2 p = __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&pw[k]));
        float2 g = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&gw[k]));
        float2 m = __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&mw[k]));
        float2 v = __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&vw[k]));
        g.x = __bfloat162float(__float2bfloat16_rn(g.x * coef));
        g.y = __bfloat162float(__float2bfloat16_rn(g.y * coef));
        adamw_one(p.x, g.x, m.x, v.x, a);
        adamw_one(p.y, g.y, m.y, v.y, a);
        __nv_bfloat162 pb = __floats2bfloat162_rn(p.x, p.y), mb = __floats2bfloat162_rn(m.x, m.y),
                       vb = __floats2bfloat162_rn(v.x, v.y);
        pw[k] = *reinterpret_cast<uint32_t*>(&pb);
        mw[k] = *reinterpret_cast<uint32_t*>(&mb);
        vw[k] = *reinterpret_cast<uint32_t*>(&vb);
      }
      reinterpret_cast<uint4*>(P)[i8] = p4;
      reinterpret_cast<uint4*>(M)[i8] = m4;
      reinterpret_cast<uint4*>(V)[i8] = v4;
    }
  } else {
    for (size_t i = start + threadIdx.x; i < end; i += MT_THREADS) {
      float p = __bfloat162float(P[i]), m = __bfloat162float(M[i]), v = __bfloat162float(V[i]);
      const float g = __bfloat162float(__float2bfloat16_rn(__bfloat162float(G[i]) * coef));
      adamw_one(p, g, m, v, a);
      P[i] = __float2bfloat16_rn(p);
      M[i] = __float2bfloat16_rn(m);
      V[i] = __float2bfloat16_rn(v);
    }
  }
}

}  // namespace b200

using namespace b200;

namespace {

long long total_blocks(const size_t* numels, int n) {
  long long b = 0;
  for (int i = 0; i < n; ++i) b += static_cast<long long>((numels[i] + MT_CHUNK - 1) / MT_CHUNK);
  return b;
}

// Fills a table for tensors [first, first+count); returns its block count.
int fill_table(TensorTable& t, void* const* p, void* const* g, void* const* m, void* const* v, const size_t* numels,
               int first, int count) {
  t.n = count;
  int blocks = 0;
  for (int i = 0; i < count; ++i) {
    t.p[i] = p ? p[first + i] : nullptr;
    t.g[i] = g[first + i];
    t.m[i] = m ? m[first + i] : nullptr;
    t.v[i] = v ? v[first + i] : nullptr;
    t.numel[i] = numels[first + i];
    t.first_block[i] = blocks;
    blocks += static_cast<int>((numels[first + i] + MT_CHUNK - 1) / MT_CHUNK);
  }
  t.first_block[count] = blocks;
  return blocks;
}

}  // namespace

extern "C" {

size_t b200sft_grad_norm_workspace_bytes(const size_t* numels, int n_tensors) {
  if (!numels || n_tensors <= 0) return 0;
  return static_cast<size_t>(total_blocks(numels, n_tensors)) * sizeof(double);
}

int b200sft_grad_norm_bf16(void* const* grads, const size_t* numels, int n_tensors, float* total_norm, void* workspace,
                           size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (!grads || !numels || n_tensors <= 0 || !total_norm) return fail(B200SFT_ERR_INVALID, "grad_norm: bad arguments");
  const long long blocks = total_blocks(numels, n_tensors);
  if (blocks > 0x7fffffffLL) return fail(B200SFT_ERR_UNSUPPORTED, "grad_norm: too many elements");
  if (workspace_bytes < static_cast<size_t>(blocks) * sizeof(double) || (blocks > 0 && !workspace))
    return fail(B200SFT_ERR_WORKSPACE, "grad_norm workspace too small: have %zu, need %zu", workspace_bytes,
                static_cast<size_t>(blocks) * sizeof(double));
  double* partial = static_cast<double*>(workspace);
  int done_blocks = 0;
  for (int first = 0; first < n_tensors; first += MT_MAX_TENSORS) {
    const int count = n_tensors - first < MT_MAX_TENSORS ? n_tensors - first : MT_MAX_TENSORS;
    TensorTable t;
    const int nb = fill_table(t, nullptr, grads, nullptr, nullptr, numels, first, count);
    if (nb > 0) {
      grad_sumsq_kernel<<<nb, MT_THREADS, 0, stream>>>(t, done_blocks, partial);
      B200_CUDA_CHECK(cudaGetLastError());
    }
    done_blocks += nb;
  }
  grad_norm_finish_kernel<<<1, 1024, 0, stream>>>(partial, done_blocks, total_norm);
  B200_CUDA_CHECK(cudaGetLastError());
  return B200SFT_OK;
}

int b200sft_adamw_clip_bf16(void* const* params, void* const* grads, void* const* exp_avgs, void* const* exp_avg_sqs,
                            const size_t* numels, int n_tensors, float lr, float beta1, float beta2, float eps,
                            float weight_decay, int step, float max_norm, const float* total_norm, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (!params || !grads || !exp_avgs || !exp_avg_sqs || !numels || n_tensors <= 0)
    return fail(B200SFT_ERR_INVALID, "adamw: bad arguments");
  if (step < 1) return fail(B200SFT_ERR_INVALID, "adamw: step counts from 1 (got %d)", step);
  if (max_norm > 0.f && !total_norm) return fail(B200SFT_ERR_INVALID, "adamw: clipping needs the total_norm scalar");
  AdamArgs a;
  a.lr = lr;
  a.beta1 = beta1;
  a.beta2 = beta2;
  a.eps = eps;
  a.weight_decay = weight_decay;
  a.bias1 = static_cast<float>(1.0 - pow(static_cast<double>(beta1), step));
  a.bias2_sqrt = static_cast<float>(sqrt(1.0 - pow(static_cast<double>(beta2), step)));
  a.max_norm = max_norm;
  for (int first = 0; first < n_tensors; first += MT_MAX_TENSORS) {
    const int count = n_tensors - first < MT_MAX_TENSORS ? n_tensors - first : MT_MAX_TENSORS;
    TensorTable t;
    const int nb = fill_table(t, params, grads, exp_avgs, exp_avg_sqs, numels, first, count);
    if (nb > 0) {
      adamw_clip_bf16_kernel<<<nb, MT_THREADS, 0, stream>>>(t, a, total_norm);
      B200_CUDA_CHECK(cudaGetLastError());
    }
  }
  return B200SFT_OK;
}

}  // extern "C"
