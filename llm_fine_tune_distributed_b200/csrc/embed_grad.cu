// Tied-embedding gradient, exchanged as rows instead of as a dense [V,H] tensor.
//
// lm_head.weight is embed_tokens.weight (SP/transformers/models/smollm3/modeling_smollm3.py:367,447) and the
// reference trains it (training.py:136-141).  Its gradient has two parts: the head's dense dW, final at the very
// start of backward(), and the embedding's scatter-add of dY rows (the gradient w.r.t. the embedding output), the
// very last op of backward().  In stock DDP the second part keeps the 521 MiB bucket of that parameter from being
// reduced until backward() has ended.  Averaging over ranks is linear, so the two parts can be averaged separately:
// the dense part goes through the bucket hook early (overlapping the trunk's backward), and the embedding part is
// what this file handles -- every rank publishes its (token ids, dY rows) in symmetric memory and then adds
//
//     grad[v, :] = bf16( grad[v, :] + scale * sum over ranks r, positions t with ids_r[t] == v of dY_r[t, :] )
//
// for the touched rows v only, reading the peers' rows straight over NVLink (plain loads from peer pointers; no
// all-gather buffer, every row crosses the link once per reader).  At most world * tokens rows of 4 KB move instead
// of 525 MB.
//
// The sum must have ONE order on every rank (replicas would drift apart otherwise) and the same order from run
// to run, so the tokens are grouped by vocabulary row with a stable counting sort -- integer atomics only for
// counts, never for values -- and added in fp32 in ascending (rank, position) order:
//   hist      per 1024-token block b: hist[b][v] = tokens of block b with id v
//   prefix    per v: exclusive prefix of hist[.][v] over blocks; count[v]
//   scan      two kernels over 1024-row tiles: start[v] = exclusive scan of count; list of touched rows; 64-token
//             chunks per row and the work-item -> row map
//   place     token j of block b goes to perm[start[v] + hist[b][v] + (tokens with the same id before j in b)]
//   sum       per (row, chunk): fp32 sum of <= 64 dY rows in perm order; a row with one chunk is finished here
//   finish    rows with several chunks: chunk partials added in order, then the same update
// Long segments (a frequent token, many ranks) are split into chunks so that no single thread block walks them alone.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/b200sft.h"
#include "capi_common.h"

namespace b200 {
namespace {

constexpr int EG_MAX_PEERS = 16;
constexpr int EG_BLOCK_TOKENS = 1024;  // tokens per histogram block (= threads of hist / place)
constexpr int EG_CHUNK = 64;           // tokens summed by one work item
constexpr int EG_SUM_THREADS = 256;
constexpr int EG_MAX_VEC = 4;          // 16-byte column vectors per thread: H <= 256 * 4 * 8 = 8192

struct EgPeers {
  const int32_t* ids[EG_MAX_PEERS];         // [cap] token ids of rank r (negative = skip: padding)
  const __nv_bfloat16* dy[EG_MAX_PEERS];    // [cap, H] gradient w.r.t. the embedding output of rank r
  const int32_t* n_tokens[EG_MAX_PEERS];    // device scalar: valid prefix length of rank r's arrays (<= cap)
};

// One element of the four simultaneous scans over the vocabulary rows: tokens, touched rows, chunks, partial slots.
struct EgQuad {
  int tok, row, wrk, par;
};

struct EgScratch {
  int32_t* hist;     // [n_blocks][V]
  int32_t* count;    // [V]
  int32_t* start;    // [V]
  int32_t* rows;     // [V]   touched rows, ascending
  int32_t* wstart;   // [V+1] first work item of touched row u (exclusive scan of its chunk count)
  int32_t* pstart;   // [V+1] first partial-sum slot of touched row u (rows with more than one chunk own slots)
  int32_t* wrow;     // [max_work] touched-row index u of every work item
  EgQuad* tile_sums; // [ceil(V / 1024)]
  int32_t* totals;   // [2]   n_rows, n_work
  int32_t* perm;     // [world * cap]
  float* partial;    // [max_partials][H]
};

__device__ __forceinline__ int eg_token_id(const EgPeers& p, int r, int t, int cap, int V) {
  if (t >= cap) return -1;
  const int n = min(max(__ldg(p.n_tokens[r]), 0), cap);
  if (t >= n) return -1;
  const int id = p.ids[r][t];
  return (id >= 0 && id < V) ? id : -1;
}

// grid = world * blocks_per_rank, block = EG_BLOCK_TOKENS
__global__ void __launch_bounds__(EG_BLOCK_TOKENS)
    eg_hist_kernel(EgPeers p, int cap, int blocks_per_rank, int V, int32_t* __restrict__ hist) {
  const int r = blockIdx.x / blocks_per_rank, c = blockIdx.x % blocks_per_rank;
  const int id = eg_token_id(p, r, c * EG_BLOCK_TOKENS + threadIdx.x, cap, V);
  if (id >= 0) atomicAdd(&hist[static_cast<size_t>(blockIdx.x) * V + id], 1);  // integer: the result has no order
}

__global__ void eg_prefix_blocks_kernel(int32_t* __restrict__ hist, int n_blocks, int V, int32_t* __restrict__ count) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= V) return;
  int run = 0;
  for (int b0 = 0; b0 < n_blocks; b0 += 8) {  // eight independent loads in flight, then the running sum
    int c[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) c[i] = (b0 + i < n_blocks) ? hist[static_cast<size_t>(b0 + i) * V + v] : 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (b0 + i < n_blocks) hist[static_cast<size_t>(b0 + i) * V + v] = run;
      run += c[i];
    }
  }
  count[v] = run;
}

__device__ __forceinline__ int eg_chunks(int c) { return (c + EG_CHUNK - 1) / EG_CHUNK; }
__device__ __forceinline__ int eg_slots(int c) { return c > EG_CHUNK ? eg_chunks(c) : 0; }

// Four exclusive scans over the V counts (tokens, touched rows, 64-token chunks, partial-sum slots), in two parallel
// kernels over tiles of 1024 vocabulary rows: tile sums first, then every tile adds up the sums of the tiles before it
// (at most a few hundred values) and scans its own rows.
__device__ __forceinline__ EgQuad eg_quad(int c) { return EgQuad{c, c > 0 ? 1 : 0, eg_chunks(c), eg_slots(c)}; }
__device__ __forceinline__ EgQuad operator+(EgQuad a, EgQuad b) {
  return EgQuad{a.tok + b.tok, a.row + b.row, a.wrk + b.wrk, a.par + b.par};
}
__device__ __forceinline__ EgQuad eg_shfl_up(EgQuad q, int d) {
  return EgQuad{__shfl_up_sync(0xffffffffu, q.tok, d), __shfl_up_sync(0xffffffffu, q.row, d),
                __shfl_up_sync(0xffffffffu, q.wrk, d), __shfl_up_sync(0xffffffffu, q.par, d)};
}
// Inclusive scan across the 1024 threads of a block; returns this thread's inclusive value, *total = block sum.
__device__ __forceinline__ EgQuad eg_block_scan(EgQuad q, EgQuad* s_warp /*[32]*/, EgQuad* total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int d = 1; d < 32; d <<= 1) {
    const EgQuad o = eg_shfl_up(q, d);
    if (lane >= d) q = q + o;
  }
  if (lane == 31) s_warp[warp] = q;
  __syncthreads();
  if (warp == 0) {
    EgQuad w = s_warp[lane];
    for (int d = 1; d < 32; d <<= 1) {
      const EgQuad o = eg_shfl_up(w, d);
      if (lane >= d) w = w + o;
    }
    s_warp[lane] = w;
  }
  __syncthreads();
  if (warp > 0) q = q + s_warp[warp - 1];
  *total = s_warp[31];
  return q;
}

__global__ void __launch_bounds__(1024)
    eg_tile_sums_kernel(const int32_t* __restrict__ count, int V, EgQuad* __restrict__ tile_sums) {
  __shared__ EgQuad s_warp[32];
  const int v = blockIdx.x * 1024 + threadIdx.x;
  EgQuad total;
  eg_block_scan(eg_quad(v < V ? count[v] : 0), s_warp, &total);
  if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}

__global__ void __launch_bounds__(1024)
    eg_scan_rows_kernel(const int32_t* __restrict__ count, int V, const EgQuad* __restrict__ tile_sums,
                        int32_t* __restrict__ start, int32_t* __restrict__ rows, int32_t* __restrict__ wstart,
                        int32_t* __restrict__ pstart, int32_t* __restrict__ wrow, int32_t* __restrict__ totals) {
  __shared__ EgQuad s_warp[32];
  __shared__ EgQuad s_base;
  if (threadIdx.x < 32) {  // sum of the tiles before this one (fixed order; integers)
    EgQuad b = {0, 0, 0, 0};
    for (int t = threadIdx.x; t < static_cast<int>(blockIdx.x); t += 32) b = b + tile_sums[t];
    for (int d = 16; d > 0; d >>= 1)
      b = b + EgQuad{__shfl_down_sync(0xffffffffu, b.tok, d), __shfl_down_sync(0xffffffffu, b.row, d),
                     __shfl_down_sync(0xffffffffu, b.wrk, d), __shfl_down_sync(0xffffffffu, b.par, d)};
    if (threadIdx.x == 0) s_base = b;
  }
  __syncthreads();
  const int v = blockIdx.x * 1024 + threadIdx.x;
  const int c = v < V ? count[v] : 0;
  const EgQuad mine = eg_quad(c);
  EgQuad total;
  const EgQuad incl = eg_block_scan(mine, s_warp, &total);
  const EgQuad base = s_base;
  const int tok = base.tok + incl.tok - mine.tok, row = base.row + incl.row - mine.row;
  const int wrk = base.wrk + incl.wrk - mine.wrk, par = base.par + incl.par - mine.par;
  if (v < V) start[v] = tok;
  if (c > 0) {
    rows[row] = v;
    wstart[row] = wrk;
    pstart[row] = par;
    for (int k = 0; k < mine.wrk; ++k) wrow[wrk + k] = row;  // work item -> touched row (long only for very frequent tokens)
  }
  if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 1023) {
    const EgQuad all = base + total;
    totals[0] = all.row;
    totals[1] = all.wrk;
    wstart[all.row] = all.wrk;  // sentinels: end of the last row's work items / slots
    pstart[all.row] = all.par;
  }
}

__global__ void __launch_bounds__(EG_BLOCK_TOKENS)
    eg_place_kernel(EgPeers p, int cap, int blocks_per_rank, int V, const int32_t* __restrict__ hist,
                    const int32_t* __restrict__ start, int32_t* __restrict__ perm) {
  __shared__ int s_id[EG_BLOCK_TOKENS];
  const int r = blockIdx.x / blocks_per_rank, c = blockIdx.x % blocks_per_rank;
  const int t = c * EG_BLOCK_TOKENS + threadIdx.x;
  const int id = eg_token_id(p, r, t, cap, V);
  s_id[threadIdx.x] = id;
  __syncthreads();
  if (id < 0) return;
  int before = 0;  // same id earlier in this block: keeps the sort stable without any ordering between threads
  for (int i = 0; i < static_cast<int>(threadIdx.x); ++i) before += (s_id[i] == id);
  perm[start[id] + hist[static_cast<size_t>(blockIdx.x) * V + id] + before] = r * cap + t;
}

__device__ __forceinline__ void eg_add8(float (&a)[8], uint4 v) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const float2 f = __bfloat1622float2(h[q]);
    a[2 * q] += f.x;
    a[2 * q + 1] += f.y;
  }
}

// grad[row, 8 columns] = bf16(grad + scale * a)
__device__ __forceinline__ void eg_update8(__nv_bfloat16* g, const float (&a)[8], float scale) {
  uint4 old = *reinterpret_cast<uint4*>(g);
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&old);
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const float2 f = __bfloat1622float2(h[q]);
    h[q] = __floats2bfloat162_rn(__fmaf_rn(scale, a[2 * q], f.x), __fmaf_rn(scale, a[2 * q + 1], f.y));
  }
  *reinterpret_cast<uint4*>(g) = old;
}

// One work item = (touched row u, chunk k of its token segment) per block iteration.
__global__ void __launch_bounds__(EG_SUM_THREADS)
    eg_sum_chunks_kernel(EgPeers p, int cap, int H, float scale, const int32_t* __restrict__ count,
                         const int32_t* __restrict__ start, const int32_t* __restrict__ rows,
                         const int32_t* __restrict__ wstart, const int32_t* __restrict__ pstart,
                         const int32_t* __restrict__ wrow, const int32_t* __restrict__ totals,
                         const int32_t* __restrict__ perm, float* __restrict__ partial, __nv_bfloat16* __restrict__ grad) {
  const int n_work = totals[1];
  const int nvec = H / 8;
  __shared__ int s_tok[EG_CHUNK];
  for (int w = blockIdx.x; w < n_work; w += gridDim.x) {
    const int u = wrow[w], row = rows[u], k = w - wstart[u];
    const int L = count[row], first = k * EG_CHUNK, len = min(EG_CHUNK, L - first);
    __syncthreads();  // s_tok of the previous work item is no longer read
    if (threadIdx.x < len) s_tok[threadIdx.x] = perm[start[row] + first + threadIdx.x];
    __syncthreads();
    float acc[EG_MAX_VEC][8];
#pragma unroll
    for (int i = 0; i < EG_MAX_VEC; ++i)
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[i][e] = 0.f;
    for (int j = 0; j < len; ++j) {  // ascending (rank, position): the one summation order
      const int tok = s_tok[j], r = tok / cap, t = tok - r * cap;
      const uint4* src = reinterpret_cast<const uint4*>(p.dy[r] + static_cast<size_t>(t) * H);
#pragma unroll
      for (int i = 0; i < EG_MAX_VEC; ++i) {
        const int v = threadIdx.x + i * EG_SUM_THREADS;
        if (v < nvec) eg_add8(acc[i], src[v]);
      }
    }
    const bool whole_row = (L <= EG_CHUNK);
#pragma unroll
    for (int i = 0; i < EG_MAX_VEC; ++i) {
      const int v = threadIdx.x + i * EG_SUM_THREADS;
      if (v >= nvec) continue;
      if (whole_row) {
        eg_update8(grad + static_cast<size_t>(row) * H + v * 8, acc[i], scale);
      } else {
        float4* dst = reinterpret_cast<float4*>(partial + static_cast<size_t>(pstart[u] + k) * H + v * 8);
        dst[0] = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
        dst[1] = make_float4(acc[i][4], acc[i][5], acc[i][6], acc[i][7]);
      }
    }
  }
}

__global__ void __launch_bounds__(EG_SUM_THREADS)
    eg_finish_rows_kernel(int H, float scale, const int32_t* __restrict__ count, const int32_t* __restrict__ rows,
                          const int32_t* __restrict__ pstart, const int32_t* __restrict__ totals,
                          const float* __restrict__ partial, __nv_bfloat16* __restrict__ grad) {
  const int n_rows = totals[0];
  const int nvec = H / 8;
  for (int u = blockIdx.x; u < n_rows; u += gridDim.x) {
    const int row = rows[u];
    if (count[row] <= EG_CHUNK) continue;  // finished by the chunk kernel
    const int w0 = pstart[u], w1 = pstart[u + 1];
    for (int v = threadIdx.x; v < nvec; v += EG_SUM_THREADS) {
      float a[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      for (int w = w0; w < w1; ++w) {  // chunk order = token order
        const float4* src = reinterpret_cast<const float4*>(partial + static_cast<size_t>(w) * H + v * 8);
        const float4 x = src[0], y = src[1];
        a[0] += x.x; a[1] += x.y; a[2] += x.z; a[3] += x.w;
        a[4] += y.x; a[5] += y.y; a[6] += y.z; a[7] += y.w;
      }
      eg_update8(grad + static_cast<size_t>(row) * H + v * 8, a, scale);
    }
  }
}

inline size_t eg_align(size_t x) { return (x + 255) / 256 * 256; }

int eg_blocks_per_rank(int cap) { return (cap + EG_BLOCK_TOKENS - 1) / EG_BLOCK_TOKENS; }

// Work items: ceil(L / EG_CHUNK) per touched row <= tokens / EG_CHUNK + touched rows, touched rows <= min(tokens, V).
size_t eg_max_work(int world, int cap, int V) {
  const size_t tokens = static_cast<size_t>(world) * cap;
  return tokens / EG_CHUNK + (tokens < static_cast<size_t>(V) ? tokens : static_cast<size_t>(V)) + 1;
}

// Partial-sum slots: only rows with more than EG_CHUNK tokens own any, ceil(L / EG_CHUNK) <= L / EG_CHUNK + 1 each
// and there are at most tokens / (EG_CHUNK + 1) such rows.
size_t eg_max_partials(int world, int cap) {
  const size_t tokens = static_cast<size_t>(world) * cap;
  return tokens / EG_CHUNK + tokens / (EG_CHUNK + 1) + 1;
}

size_t eg_workspace(int world, int cap, int V, int H, EgScratch* s, void* base) {
  uint8_t* p = static_cast<uint8_t*>(base);
  size_t off = 0;
  auto take = [&](size_t bytes) {
    uint8_t* at = p ? p + off : nullptr;
    off += eg_align(bytes);
    return at;
  };
  const size_t nb = static_cast<size_t>(world) * eg_blocks_per_rank(cap);
  int32_t* hist = reinterpret_cast<int32_t*>(take(nb * V * sizeof(int32_t)));
  int32_t* count = reinterpret_cast<int32_t*>(take(static_cast<size_t>(V) * 4));
  int32_t* start = reinterpret_cast<int32_t*>(take(static_cast<size_t>(V) * 4));
  int32_t* rows = reinterpret_cast<int32_t*>(take(static_cast<size_t>(V) * 4));
  int32_t* wstart = reinterpret_cast<int32_t*>(take((static_cast<size_t>(V) + 1) * 4));
  int32_t* pstart = reinterpret_cast<int32_t*>(take((static_cast<size_t>(V) + 1) * 4));
  int32_t* wrow = reinterpret_cast<int32_t*>(take(eg_max_work(world, cap, V) * 4));
  EgQuad* tile_sums = reinterpret_cast<EgQuad*>(take(static_cast<size_t>((V + 1023) / 1024) * sizeof(EgQuad)));
  int32_t* totals = reinterpret_cast<int32_t*>(take(16));
  int32_t* perm = reinterpret_cast<int32_t*>(take(static_cast<size_t>(world) * cap * 4));
  float* partial = reinterpret_cast<float*>(take(eg_max_partials(world, cap) * H * sizeof(float)));
  if (s) *s = EgScratch{hist, count, start, rows, wstart, pstart, wrow, tile_sums, totals, perm, partial};
  return off;
}

}  // namespace
}  // namespace b200

using namespace b200;

extern "C" {

size_t b200sft_embed_grad_workspace_bytes(int world, int cap, int V, int H) {
  if (world <= 0 || cap <= 0 || V <= 0 || H <= 0) return 0;
  return eg_workspace(world, cap, V, H, nullptr, nullptr);
}

int b200sft_embed_grad_rows(const uint64_t* ids_ptrs_host, const uint64_t* dy_ptrs_host, const uint64_t* n_ptrs_host,
                            int world, int cap, int V, int H, float scale, void* grad, void* workspace,
                            size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (!ids_ptrs_host || !dy_ptrs_host || !n_ptrs_host || !grad) return fail(B200SFT_ERR_INVALID, "null pointer");
  if (world <= 0 || world > EG_MAX_PEERS) return fail(B200SFT_ERR_INVALID, "world %d out of range (1..%d)", world, EG_MAX_PEERS);
  if (cap <= 0 || V <= 0 || H <= 0) return fail(B200SFT_ERR_INVALID, "cap, V, H must be positive (got %d, %d, %d)", cap, V, H);
  if (H % 8 != 0 || H > EG_SUM_THREADS * EG_MAX_VEC * 8)
    return fail(B200SFT_ERR_INVALID, "H must be a multiple of 8 and at most %d (got %d)", EG_SUM_THREADS * EG_MAX_VEC * 8, H);
  if (static_cast<long long>(world) * cap >= (1LL << 31)) return fail(B200SFT_ERR_INVALID, "world * cap must stay below 2^31");
  if ((reinterpret_cast<uintptr_t>(grad) & 15) != 0) return fail(B200SFT_ERR_INVALID, "grad must be 16-byte aligned");
  const size_t need = b200sft_embed_grad_workspace_bytes(world, cap, V, H);
  if (workspace == nullptr || workspace_bytes < need)
    return fail(B200SFT_ERR_WORKSPACE, "embed_grad_rows workspace too small: have %zu, need %zu", workspace_bytes, need);
  EgPeers p = {};
  for (int r = 0; r < world; ++r) {
    if (!ids_ptrs_host[r] || !dy_ptrs_host[r] || !n_ptrs_host[r] || (dy_ptrs_host[r] & 15) != 0)
      return fail(B200SFT_ERR_INVALID, "bad pointer for rank %d (dY must be 16-byte aligned)", r);
    p.ids[r] = reinterpret_cast<const int32_t*>(ids_ptrs_host[r]);
    p.dy[r] = reinterpret_cast<const __nv_bfloat16*>(dy_ptrs_host[r]);
    p.n_tokens[r] = reinterpret_cast<const int32_t*>(n_ptrs_host[r]);
  }
  EgScratch s;
  eg_workspace(world, cap, V, H, &s, workspace);
  const int bpr = eg_blocks_per_rank(cap), nb = world * bpr;
  B200_CUDA_CHECK(cudaMemsetAsync(s.hist, 0, static_cast<size_t>(nb) * V * sizeof(int32_t), stream));
  eg_hist_kernel<<<nb, EG_BLOCK_TOKENS, 0, stream>>>(p, cap, bpr, V, s.hist);
  B200_CUDA_CHECK(cudaGetLastError());
  eg_prefix_blocks_kernel<<<(V + 255) / 256, 256, 0, stream>>>(s.hist, nb, V, s.count);
  B200_CUDA_CHECK(cudaGetLastError());
  const int v_tiles = (V + 1023) / 1024;
  eg_tile_sums_kernel<<<v_tiles, 1024, 0, stream>>>(s.count, V, s.tile_sums);
  B200_CUDA_CHECK(cudaGetLastError());
  eg_scan_rows_kernel<<<v_tiles, 1024, 0, stream>>>(s.count, V, s.tile_sums, s.start, s.rows, s.wstart, s.pstart, s.wrow,
                                                    s.totals);
  B200_CUDA_CHECK(cudaGetLastError());
  eg_place_kernel<<<nb, EG_BLOCK_TOKENS, 0, stream>>>(p, cap, bpr, V, s.hist, s.start, s.perm);
  B200_CUDA_CHECK(cudaGetLastError());
  int dev = 0, sms = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (sms <= 0) sms = 148;
  eg_sum_chunks_kernel<<<sms * 8, EG_SUM_THREADS, 0, stream>>>(p, cap, H, scale, s.count, s.start, s.rows, s.wstart, s.pstart,
                                                               s.wrow, s.totals, s.perm, s.partial,
                                                               static_cast<__nv_bfloat16*>(grad));
  B200_CUDA_CHECK(cudaGetLastError());
  eg_finish_rows_kernel<<<sms * 4, EG_SUM_THREADS, 0, stream>>>(H, scale, s.count, s.rows, s.pstart, s.totals, s.partial,
                                                                static_cast<__nv_bfloat16*>(grad));
  B200_CUDA_CHECK(cudaGetLastError());
  return B200SFT_OK;
}

}  // extern "C"
