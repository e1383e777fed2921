// HBM-bound integer / byte kernels of the SFT hot path: the label front-end (shift + counts) and
// the DDP gradient-bucket scale / cast kernels.  The streaming kernels give each block one
// contiguous 16 KiB run (measured faster on B200 than a persistent grid-stride sweep, whose
// per-thread accesses land megabytes apart); vector accesses are 8/16 bytes, 4 in flight per thread.
#include <cuda_bf16.h>
#include <cooperative_groups.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/b200sft.h"
#include "capi_common.h"

namespace cg = cooperative_groups;

namespace b200 {

char* last_error_buf() {
  static thread_local char buf[512] = {0};
  return buf;
}

namespace {

int sm_count() {
  int dev = 0, n = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
  return n > 0 ? n : 148;
}

int grid_for(size_t work_items, int threads, int ctas_per_sm) {
  size_t blocks = (work_items + threads - 1) / threads;
  size_t cap = static_cast<size_t>(sm_count()) * ctas_per_sm;
  if (blocks > cap) blocks = cap;
  if (blocks == 0) blocks = 1;
  return static_cast<int>(blocks);
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// label front-end
// ---------------------------------------------------------------------------------------------
// shifted[b,t] = labels[b,t+1] (t < T-1), ignore_index at t = T-1; counts of non-ignored entries
// before and after the shift.  Integer-exact; counts accumulate with 64-bit atomics (order-free).
__global__ void label_prep_kernel(const int64_t* __restrict__ labels, int B, int T, int64_t ignore_index,
                                  int32_t* __restrict__ shifted, unsigned long long* __restrict__ counts) {
  const size_t total = static_cast<size_t>(B) * T;
  unsigned int c_unshifted = 0, c_shifted = 0;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int t = static_cast<int>(i % T);
    const int64_t cur = labels[i];
    c_unshifted += (cur != ignore_index) ? 1u : 0u;
    int64_t nxt = ignore_index;
    if (t + 1 < T) nxt = labels[i + 1];
    c_shifted += (nxt != ignore_index) ? 1u : 0u;
    shifted[i] = static_cast<int32_t>(nxt);
  }
  // warp -> block -> one atomic pair per block
  for (int off = 16; off > 0; off >>= 1) {
    c_unshifted += __shfl_down_sync(0xffffffffu, c_unshifted, off);
    c_shifted += __shfl_down_sync(0xffffffffu, c_shifted, off);
  }
  __shared__ unsigned int s_u[32], s_s[32];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) {
    s_u[warp] = c_unshifted;
    s_s[warp] = c_shifted;
  }
  __syncthreads();
  if (warp == 0) {
    const int nw = (blockDim.x + 31) >> 5;
    unsigned int u = lane < nw ? s_u[lane] : 0u, s = lane < nw ? s_s[lane] : 0u;
    for (int off = 16; off > 0; off >>= 1) {
      u += __shfl_down_sync(0xffffffffu, u, off);
      s += __shfl_down_sync(0xffffffffu, s, off);
    }
    if (lane == 0) {
      atomicAdd(&counts[0], static_cast<unsigned long long>(u));
      atomicAdd(&counts[1], static_cast<unsigned long long>(s));
    }
  }
}

// ---------------------------------------------------------------------------------------------
// label front-end, whole: collator masks + shift + both counts + live-row compaction, ONE launch
// ---------------------------------------------------------------------------------------------
// The unshifted label of token i is what the collators would have written on the host:
//   src[i] (labels, or input_ids when the collator copies them), replaced by ignore_index where
//   attention_mask[i] == 0 (padding), completion_mask[i] == 0 (prompt tokens under
//   completion-only loss) or position_ids[i] == 0 (first token of a packed sequence -- the
//   separator DataCollatorWithFlattening writes, SP/transformers/data/data_collator.py:1426-1428).
// 1024-thread blocks: each warp owns one contiguous run and walks it 32 tokens at a time
// (coalesced), so the stable compaction is a ballot + popc per step and one 32-entry scan over
// the warp totals.  Nothing needs zeroing beforehand and nothing is accumulated with atomics.
// One block up to 16 Ki tokens; beyond that one 8-CTA cluster (see the kernel).
struct LabelSources {
  const int32_t* pre_shifted;  // non-null: labels are shifted int32 already; only the compaction runs
  const int64_t* src;
  const int64_t* attention_mask;
  const int64_t* completion_mask;
  const int64_t* position_ids;
  int64_t ignore_index;
  __device__ __forceinline__ int64_t at(size_t i) const {
    int64_t v = src[i];
    if (attention_mask != nullptr && attention_mask[i] == 0) v = ignore_index;
    if (completion_mask != nullptr && completion_mask[i] == 0) v = ignore_index;
    if (position_ids != nullptr && position_ids[i] == 0) v = ignore_index;
    return v;
  }
};

constexpr int FRONTEND_THREADS = 1024;
constexpr int FRONTEND_CLUSTER = 8;               // CTAs of the cluster variant (portable maximum)
constexpr size_t FRONTEND_CLUSTER_MIN = 16384;    // tokens from which one block's latency shows (measured)

// CLUSTER == 1: one block.  CLUSTER == 8: one thread-block cluster; every CTA scans its own 32 warp
// totals, the per-CTA totals are exchanged through distributed shared memory (8 remote reads per
// CTA), so the cross-CTA prefix needs neither a second launch nor global-memory flags.
template <int CLUSTER>
__global__ void __launch_bounds__(FRONTEND_THREADS)
label_frontend_kernel(LabelSources in, int B, int T, int64_t* __restrict__ labels_out, int32_t* __restrict__ shifted,
                      int64_t* __restrict__ counts, int32_t* __restrict__ row_index, int32_t* __restrict__ labels_c,
                      int32_t* __restrict__ pos, int32_t* __restrict__ n_live) {
  __shared__ unsigned int s_live[32], s_unshifted[32], s_base[32];
  __shared__ unsigned int s_cta_total[2];  // this CTA's live / unshifted totals (read by the other CTAs)
  __shared__ unsigned int s_grand[3];      // live rows in lower-ranked CTAs, all live rows, all unshifted
  unsigned int cta_rank = 0;
  if constexpr (CLUSTER > 1) cta_rank = cg::this_cluster().block_rank();
  const size_t total = static_cast<size_t>(B) * T;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const unsigned int lt_mask = (1u << lane) - 1u;
  const size_t n_warps = static_cast<size_t>(CLUSTER) * 32;
  const size_t per = ((total + n_warps - 1) / n_warps + 31) / 32 * 32;  // tokens per warp, a multiple of 32
  const size_t lo = min((static_cast<size_t>(cta_rank) * 32 + warp) * per, total), hi = min(lo + per, total);
  const int32_t ignore32 = static_cast<int32_t>(in.ignore_index);

  // pass 1: labels, shifted labels, per-warp counts
  unsigned int live = 0, unshifted = 0;
  for (size_t base = lo; base < hi; base += 32) {
    const size_t i = base + lane;
    bool is_live = false, is_unshifted = false;
    if (i < hi) {
      if (in.pre_shifted != nullptr) {
        is_live = in.pre_shifted[i] != ignore32;
      } else {
        const int64_t cur = in.at(i);
        const int t = static_cast<int>(i % T);
        const int64_t nxt = (t + 1 < T) ? in.at(i + 1) : in.ignore_index;
        if (labels_out != nullptr) labels_out[i] = cur;
        shifted[i] = static_cast<int32_t>(nxt);
        is_unshifted = cur != in.ignore_index;
        is_live = nxt != in.ignore_index;
      }
    }
    live += __popc(__ballot_sync(0xffffffffu, is_live));
    unshifted += __popc(__ballot_sync(0xffffffffu, is_unshifted));
  }
  if (lane == 0) {
    s_live[warp] = live;
    s_unshifted[warp] = unshifted;
  }
  __syncthreads();
  if (warp == 0) {  // exclusive scan of this CTA's 32 warp totals
    unsigned int v = s_live[lane], u = s_unshifted[lane];
    unsigned int incl = v;
    for (int off = 1; off < 32; off <<= 1) {
      const unsigned int o = __shfl_up_sync(0xffffffffu, incl, off);
      if (lane >= off) incl += o;
      u += __shfl_xor_sync(0xffffffffu, u, off);
    }
    s_base[lane] = incl - v;
    if (lane == 31) s_cta_total[0] = incl;
    if (lane == 0) s_cta_total[1] = u;
  }
  if constexpr (CLUSTER > 1) {
    cg::cluster_group cluster = cg::this_cluster();
    cluster.sync();  // every CTA's totals are in its shared memory
    if (warp == 0) {
      unsigned int l = 0, u = 0;
      if (lane < CLUSTER) {
        const unsigned int* remote = cluster.map_shared_rank(s_cta_total, lane);
        l = remote[0];
        u = remote[1];
      }
      unsigned int below = lane < static_cast<int>(cta_rank) ? l : 0u;
      for (int off = 16; off > 0; off >>= 1) {
        l += __shfl_xor_sync(0xffffffffu, l, off);
        u += __shfl_xor_sync(0xffffffffu, u, off);
        below += __shfl_xor_sync(0xffffffffu, below, off);
      }
      if (lane == 0) {
        s_grand[0] = below;
        s_grand[1] = l;
        s_grand[2] = u;
      }
    }
    cluster.sync();  // s_grand is visible; no CTA leaves while a neighbour may still read its totals
  } else {
    __syncthreads();
    if (threadIdx.x == 0) {
      s_grand[0] = 0;
      s_grand[1] = s_cta_total[0];
      s_grand[2] = s_cta_total[1];
    }
    __syncthreads();
  }
  const unsigned int n_total = s_grand[1];
  if (cta_rank == 0 && threadIdx.x == 0) {
    if (counts != nullptr) {
      counts[0] = static_cast<int64_t>(s_grand[2]);
      counts[1] = static_cast<int64_t>(n_total);
    }
    if (n_live != nullptr) *n_live = static_cast<int32_t>(n_total);
  }
  if (row_index == nullptr) return;

  // pass 2: stable compaction (each thread re-reads the shifted labels it wrote itself)
  const int32_t* __restrict__ from = in.pre_shifted != nullptr ? in.pre_shifted : shifted;
  unsigned int at = s_grand[0] + s_base[warp];
  for (size_t base = lo; base < hi; base += 32) {
    const size_t i = base + lane;
    int32_t l = ignore32;
    if (i < hi) l = from[i];
    const bool is_live = (i < hi) && l != ignore32;
    const unsigned int ballot = __ballot_sync(0xffffffffu, is_live);
    if (is_live) {
      const unsigned int j = at + __popc(ballot & lt_mask);
      row_index[j] = static_cast<int32_t>(i);
      labels_c[j] = l;
      pos[i] = static_cast<int32_t>(j);
    } else if (i < hi) {
      pos[i] = -1;
    }
    at += __popc(ballot);
  }
  for (size_t j = n_total + cta_rank * FRONTEND_THREADS + threadIdx.x; j < total;
       j += static_cast<size_t>(CLUSTER) * FRONTEND_THREADS)
    labels_c[j] = ignore32;  // dead tail
}

// ---------------------------------------------------------------------------------------------
// scale / cast (gradient buckets)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float round_bf16_f32(float a) { return __bfloat162float(__float2bfloat16_rn(a)); }
__device__ __forceinline__ uint32_t scale_bf16x2(uint32_t v, float s) {
  __nv_bfloat162 h = *reinterpret_cast<__nv_bfloat162*>(&v);
  float2 f = __bfloat1622float2(h);
  __nv_bfloat162 r = __floats2bfloat162_rn(f.x * s, f.y * s);
  return *reinterpret_cast<uint32_t*>(&r);
}

constexpr int EW_UNROLL = 4;  // independent vector accesses in flight per thread
constexpr int EW_THREADS = 256;

// Streams vectors [0, nvec): per pass a block owns EW_UNROLL * blockDim.x consecutive vectors
// (16 KiB of 16-byte vectors), thread t touching t, t + blockDim, ... so every warp access is one
// contiguous 512-byte run and all EW_UNROLL loads are issued before the first store.
template <typename VecIn, typename VecOut, class Op>
__device__ __forceinline__ void stream_vectors(const VecIn* in, VecOut* out, size_t nvec, Op op) {
  const size_t chunk = static_cast<size_t>(blockDim.x) * EW_UNROLL;
  for (size_t base = blockIdx.x * chunk; base < nvec; base += static_cast<size_t>(gridDim.x) * chunk) {
    if (base + chunk <= nvec) {
      VecIn x[EW_UNROLL];
#pragma unroll
      for (int u = 0; u < EW_UNROLL; ++u) x[u] = in[base + threadIdx.x + u * blockDim.x];
#pragma unroll
      for (int u = 0; u < EW_UNROLL; ++u) out[base + threadIdx.x + u * blockDim.x] = op(x[u]);
    } else {
      for (size_t i = base + threadIdx.x; i < nvec; i += blockDim.x) out[i] = op(in[i]);
    }
  }
}

struct ScaleBf16x8 {
  float s;
  __device__ __forceinline__ uint4 operator()(uint4 x) const {
    x.x = scale_bf16x2(x.x, s);
    x.y = scale_bf16x2(x.y, s);
    x.z = scale_bf16x2(x.z, s);
    x.w = scale_bf16x2(x.w, s);
    return x;
  }
};
struct ScaleF32x4 {
  float s;
  __device__ __forceinline__ float4 operator()(float4 x) const { return make_float4(x.x * s, x.y * s, x.z * s, x.w * s); }
};
// bf16_compress_hook rounds to bf16 BEFORE it divides (default_hooks.py:74: buffer.to(bf16).div_(world)), so the
// value is rounded, scaled in fp32 and rounded again -- identical to one rounding for power-of-two scales, and
// the reference's bits for every other world size.
__device__ __forceinline__ float compress_one(float a, float s) { return round_bf16_f32(a) * s; }
struct F32x4ToBf16x4 {
  float s;
  __device__ __forceinline__ uint2 operator()(float4 a) const {
    __nv_bfloat162 p0 = __floats2bfloat162_rn(compress_one(a.x, s), compress_one(a.y, s)),
                   p1 = __floats2bfloat162_rn(compress_one(a.z, s), compress_one(a.w, s));
    return make_uint2(*reinterpret_cast<uint32_t*>(&p0), *reinterpret_cast<uint32_t*>(&p1));
  }
};
struct Bf16x4ToF32x4 {
  float s;
  __device__ __forceinline__ float4 operator()(uint2 x) const {
    const float2 a = __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&x.x));
    const float2 b = __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&x.y));
    return make_float4(a.x * s, a.y * s, b.x * s, b.y * s);
  }
};

// In place on bf16: 2 B read + 2 B written per element.  Body on 16-byte vectors (8 elements);
// scalar head/tail (block 0) for unaligned bucket views.
__global__ void __launch_bounds__(EW_THREADS) scale_bf16_kernel(__nv_bfloat16* buf, size_t n, float s) {
  const uintptr_t addr = reinterpret_cast<uintptr_t>(buf);
  size_t head = ((16 - (addr & 15)) & 15) / 2;
  if (head > n) head = n;
  const size_t nvec = (n - head) / 8;
  uint4* v = reinterpret_cast<uint4*>(buf + head);
  stream_vectors<uint4, uint4>(v, v, nvec, ScaleBf16x8{s});
  if (blockIdx.x == 0) {
    const size_t tail0 = head + nvec * 8;
    for (size_t k = threadIdx.x; k < head + (n - tail0); k += blockDim.x) {
      const size_t j = k < head ? k : tail0 + (k - head);
      buf[j] = __float2bfloat16_rn(__bfloat162float(buf[j]) * s);
    }
  }
}

__global__ void __launch_bounds__(EW_THREADS) scale_f32_kernel(float* buf, size_t n, float s) {
  const uintptr_t addr = reinterpret_cast<uintptr_t>(buf);
  size_t head = ((16 - (addr & 15)) & 15) / 4;
  if (head > n) head = n;
  const size_t nvec = (n - head) / 4;
  float4* v = reinterpret_cast<float4*>(buf + head);
  stream_vectors<float4, float4>(v, v, nvec, ScaleF32x4{s});
  if (blockIdx.x == 0) {
    const size_t tail0 = head + nvec * 4;
    for (size_t k = threadIdx.x; k < head + (n - tail0); k += blockDim.x) {
      const size_t j = k < head ? k : tail0 + (k - head);
      buf[j] *= s;
    }
  }
}

// fp32 -> bf16 with scale (compress): 4 B read + 2 B written per element; 16-byte loads, 8-byte
// stores, 4 elements per thread access.
__global__ void __launch_bounds__(EW_THREADS)
    scale_cast_f32_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, size_t n, float s,
                               int vec_ok) {
  size_t done = 0;
  if (vec_ok) {
    const size_t nvec = n / 4;
    stream_vectors<float4, uint2>(reinterpret_cast<const float4*>(src), reinterpret_cast<uint2*>(dst), nvec,
                                  F32x4ToBf16x4{s});
    done = nvec * 4;
  }
  if (vec_ok ? blockIdx.x == 0 : true) {
    const size_t stride = vec_ok ? blockDim.x : static_cast<size_t>(gridDim.x) * blockDim.x;
    const size_t first = vec_ok ? threadIdx.x : blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
    for (size_t i = done + first; i < n; i += stride) dst[i] = __float2bfloat16_rn(compress_one(src[i], s));
  }
}

// bf16 -> fp32 with scale (decompress): 2 B read + 4 B written per element; 8-byte loads, 16-byte stores.
__global__ void __launch_bounds__(EW_THREADS)
    cast_bf16_f32_kernel(const __nv_bfloat16* __restrict__ src, float* __restrict__ dst, size_t n, float s, int vec_ok) {
  size_t done = 0;
  if (vec_ok) {
    const size_t nvec = n / 4;
    stream_vectors<uint2, float4>(reinterpret_cast<const uint2*>(src), reinterpret_cast<float4*>(dst), nvec,
                                  Bf16x4ToF32x4{s});
    done = nvec * 4;
  }
  if (vec_ok ? blockIdx.x == 0 : true) {
    const size_t stride = vec_ok ? blockDim.x : static_cast<size_t>(gridDim.x) * blockDim.x;
    const size_t first = vec_ok ? threadIdx.x : blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
    for (size_t i = done + first; i < n; i += stride) dst[i] = __bfloat162float(src[i]) * s;
  }
}

// One block per EW_UNROLL * EW_THREADS vectors (a single pass; the loop in stream_vectors only
// matters beyond 2^31 blocks).
inline unsigned int stream_grid(size_t nvec) {
  const size_t chunk = static_cast<size_t>(EW_THREADS) * EW_UNROLL;
  size_t blocks = (nvec + chunk - 1) / chunk;
  if (blocks == 0) blocks = 1;
  if (blocks > 0x7fffffffull) blocks = 0x7fffffffull;
  return static_cast<unsigned int>(blocks);
}

}  // namespace b200

using namespace b200;

namespace b200 {

// One block below FRONTEND_CLUSTER_MIN tokens, one 8-CTA cluster from there.
static int launch_label_frontend(const LabelSources& in, int B, int T, int64_t* labels_out, int32_t* shifted,
                                 int64_t* counts, int32_t* row_index, int32_t* labels_c, int32_t* pos, int32_t* n_live,
                                 cudaStream_t stream) {
  const size_t total = static_cast<size_t>(B) * T;
  if (total < FRONTEND_CLUSTER_MIN) {
    label_frontend_kernel<1><<<1, FRONTEND_THREADS, 0, stream>>>(in, B, T, labels_out, shifted, counts, row_index,
                                                                  labels_c, pos, n_live);
    B200_CUDA_CHECK(cudaGetLastError());
    return B200SFT_OK;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(FRONTEND_CLUSTER);
  cfg.blockDim = dim3(FRONTEND_THREADS);
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = FRONTEND_CLUSTER;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  B200_CUDA_CHECK(cudaLaunchKernelEx(&cfg, label_frontend_kernel<FRONTEND_CLUSTER>, in, B, T, labels_out, shifted, counts,
                                     row_index, labels_c, pos, n_live));
  return B200SFT_OK;
}

// Compaction list of already-shifted int32 labels (b200sft_compact_rows): the same kernel, shift skipped.
int launch_compact_labels(const int32_t* labels, int N, int ignore_index, int32_t* row_index, int32_t* labels_c,
                          int32_t* pos, int32_t* n_live, cudaStream_t stream) {
  LabelSources in{labels, nullptr, nullptr, nullptr, nullptr, static_cast<int64_t>(ignore_index)};
  return launch_label_frontend(in, 1, N, nullptr, nullptr, nullptr, row_index, labels_c, pos, n_live, stream);
}

}  // namespace b200

extern "C" {

const char* b200sft_last_error(void) { return last_error_buf(); }
int b200sft_version(void) { return 100; }

int b200sft_label_prep(const int64_t* labels, int B, int T, int64_t ignore_index, int32_t* shifted, int64_t* counts,
                       void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (B < 0 || T < 0) return fail(B200SFT_ERR_INVALID, "negative label shape %d x %d", B, T);
  if (!counts) return fail(B200SFT_ERR_INVALID, "null counts pointer");
  B200_CUDA_CHECK(cudaMemsetAsync(counts, 0, 2 * sizeof(int64_t), stream));
  const size_t total = static_cast<size_t>(B) * T;
  if (total == 0) return B200SFT_OK;  // empty batch: both counts are zero
  if (!labels || !shifted) return fail(B200SFT_ERR_INVALID, "null label pointer");
  const int threads = 256;
  label_prep_kernel<<<grid_for(total, threads, 8), threads, 0, stream>>>(
      labels, B, T, ignore_index, shifted, reinterpret_cast<unsigned long long*>(counts));
  B200_CUDA_CHECK(cudaGetLastError());
  return B200SFT_OK;
}

int b200sft_label_frontend(const int64_t* input_ids, const int64_t* labels, const int64_t* attention_mask,
                           const int64_t* completion_mask, const int64_t* position_ids, int B, int T,
                           int64_t ignore_index, int64_t* labels_out, int32_t* shifted, int64_t* counts,
                           int32_t* row_index, int32_t* labels_c, int32_t* pos, int32_t* n_live, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (B < 0 || T < 0) return fail(B200SFT_ERR_INVALID, "negative label shape %d x %d", B, T);
  if (!counts) return fail(B200SFT_ERR_INVALID, "null counts pointer");
  const size_t total = static_cast<size_t>(B) * T;
  if (total > static_cast<size_t>(INT32_MAX)) return fail(B200SFT_ERR_INVALID, "B*T = %zu exceeds int32 row indices", total);
  const int64_t* src = labels ? labels : input_ids;
  if (total == 0) {  // empty batch: counts and live count are zero
    B200_CUDA_CHECK(cudaMemsetAsync(counts, 0, 2 * sizeof(int64_t), stream));
    if (n_live) B200_CUDA_CHECK(cudaMemsetAsync(n_live, 0, sizeof(int32_t), stream));
    return B200SFT_OK;
  }
  if (!src || !shifted) return fail(B200SFT_ERR_INVALID, "null label source or shifted output");
  const bool compact = row_index || labels_c || pos || n_live;
  if (compact && !(row_index && labels_c && pos && n_live))
    return fail(B200SFT_ERR_INVALID, "compaction outputs must be all set or all null");
  LabelSources in{nullptr, src, attention_mask, completion_mask, position_ids, ignore_index};
  return launch_label_frontend(in, B, T, labels_out, shifted, counts, row_index, labels_c, pos, n_live, stream);
}

int b200sft_compact_labels(const int32_t* labels, int N, int ignore_index, int32_t* row_index, int32_t* labels_c,
                           int32_t* pos, int32_t* n_live, void* stream_) {
  if (N <= 0) return fail(B200SFT_ERR_INVALID, "compact_labels: bad row count %d", N);
  if (!labels || !row_index || !labels_c || !pos || !n_live) return fail(B200SFT_ERR_INVALID, "null tensor pointer");
  return launch_compact_labels(labels, N, ignore_index, row_index, labels_c, pos, n_live,
                               static_cast<cudaStream_t>(stream_));
}

int b200sft_scale_bf16_inplace(void* buf, size_t n, float scale, void* stream_) {
  if (n == 0) return B200SFT_OK;
  if (!buf) return fail(B200SFT_ERR_INVALID, "null buffer");
  if (reinterpret_cast<uintptr_t>(buf) & 1) return fail(B200SFT_ERR_INVALID, "bf16 buffer must be 2-byte aligned");
  scale_bf16_kernel<<<stream_grid((n + 7) / 8), EW_THREADS, 0, static_cast<cudaStream_t>(stream_)>>>(
      static_cast<__nv_bfloat16*>(buf), n, scale);
  B200_CUDA_CHECK(cudaGetLastError());
  return B200SFT_OK;
}

int b200sft_scale_f32_inplace(float* buf, size_t n, float scale, void* stream_) {
  if (n == 0) return B200SFT_OK;
  if (!buf) return fail(B200SFT_ERR_INVALID, "null buffer");
  if (reinterpret_cast<uintptr_t>(buf) & 3) return fail(B200SFT_ERR_INVALID, "fp32 buffer must be 4-byte aligned");
  scale_f32_kernel<<<stream_grid((n + 3) / 4), EW_THREADS, 0, static_cast<cudaStream_t>(stream_)>>>(buf, n, scale);
  B200_CUDA_CHECK(cudaGetLastError());
  return B200SFT_OK;
}

int b200sft_scale_cast_f32_to_bf16(const float* src, void* dst, size_t n, float scale, void* stream_) {
  if (n == 0) return B200SFT_OK;
  if (!src || !dst) return fail(B200SFT_ERR_INVALID, "null buffer");
  const int vec_ok = ((reinterpret_cast<uintptr_t>(src) & 15) == 0 && (reinterpret_cast<uintptr_t>(dst) & 15) == 0) ? 1 : 0;
  scale_cast_f32_bf16_kernel<<<stream_grid((n + 3) / 4), EW_THREADS, 0, static_cast<cudaStream_t>(stream_)>>>(
      src, static_cast<__nv_bfloat16*>(dst), n, scale, vec_ok);
  B200_CUDA_CHECK(cudaGetLastError());
  return B200SFT_OK;
}

int b200sft_cast_bf16_to_f32(const void* src, float* dst, size_t n, float scale, void* stream_) {
  if (n == 0) return B200SFT_OK;
  if (!src || !dst) return fail(B200SFT_ERR_INVALID, "null buffer");
  const int vec_ok = ((reinterpret_cast<uintptr_t>(src) & 15) == 0 && (reinterpret_cast<uintptr_t>(dst) & 15) == 0) ? 1 : 0;
  cast_bf16_f32_kernel<<<stream_grid((n + 3) / 4), EW_THREADS, 0, static_cast<cudaStream_t>(stream_)>>>(
      static_cast<const __nv_bfloat16*>(src), dst, n, scale, vec_ok);
  B200_CUDA_CHECK(cudaGetLastError());
  return B200SFT_OK;
}

}  // extern "C"
