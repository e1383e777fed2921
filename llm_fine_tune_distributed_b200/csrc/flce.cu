// Fused LM-head + masked cross-entropy for sm_100a: the epilogues that specialise the tcgen05
// GEMM engine (gemm_engine.cuh), the small row-wise kernels around them, and the C-ABI entry
// points declared in include/b200sft.h.
//
// Forward:  logits tile = X[128 rows] . W[256 vocab rows]^T accumulates in TMEM; the epilogue
//           folds it into a per-row (max, sum-exp) partial and picks the target logit.  A row-wise
//           kernel merges the per-vocab-tile partials into lse[N]; one block sums the loss.
// Backward: per slab of vocabulary columns, the same GEMM is recomputed and its epilogue emits
//           dlogits = (exp(z - lse) - onehot) * coef as bf16 into an [N, slab] scratch; two more
//           engine GEMMs produce dX (+)= dlogits . W[slab] and dW[slab] = dlogits^T . X, reading W,
//           X and the scratch through MN-major descriptors (no transposes in HBM).  Slabbing V
//           rather than N makes every dW slab final after one GEMM; only the small dX [N,H]
//           accumulates (fp32) across slabs.
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>

#include "../../include/b200sft.h"
#include "capi_common.h"
#include "gemm_engine.cuh"

namespace b200 {

constexpr float LOG2E = 1.4426950408889634f;

// =============================================================================================
// Epilogues
// =============================================================================================

// ---- forward: per-row online softmax statistics over one 256-wide vocabulary tile ------------
// STATS additionally carries sum exp(z - max) * z, from which the row entropy follows
// (H = lse - E_p[z]); that is what lets token-accuracy / entropy logging work without logits.
template <bool STATS>
struct FwdSoftmaxEpi {
  struct Params {
    const int32_t* labels;  // [rows]
    float2* partials;       // [tiles_n][row_pitch]  (max, sum exp(z - max))
    float* partials_ez;     // [tiles_n][row_pitch]  sum exp(z - max) * z   (STATS only)
    float* tgt_logit;       // [rows]
    int row_pitch;          // row capacity of the scratch arrays (>= live rows)
    int round_bf16;
  };
  const Params& p;
  const int M, N;
  float run_m, run_s, run_ez;
  int lbl, row;

  __device__ FwdSoftmaxEpi(const Params& p_, const GemmShape& s, uint8_t*) : p(p_), M(s.M), N(s.N) {}
  __device__ __forceinline__ void finish() {}

  __device__ __forceinline__ void tile_begin(const TileCoord& tc, int row_in_tile) {
    row = tc.m0 + row_in_tile;
    run_m = -INFINITY;
    run_s = 0.f;
    run_ez = 0.f;
    lbl = (row < M) ? __ldg(p.labels + row) : -1;
  }

  __device__ __forceinline__ void chunk(const TileCoord& tc, int, int c, const uint32_t (&acc)[32]) {
    const int col0 = tc.n0 + c;
    if (col0 >= N) return;
    float z[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) z[j] = __uint_as_float(acc[j]);
    if (p.round_bf16) {
#pragma unroll
      for (int j = 0; j < 32; ++j) z[j] = round_to_bf16(z[j]);
    }
    if (col0 + 32 > N) {  // ragged last vocabulary tile
#pragma unroll
      for (int j = 0; j < 32; ++j) z[j] = (col0 + j < N) ? z[j] : -INFINITY;
    }
    float cmax = z[0];
#pragma unroll
    for (int j = 1; j < 32; ++j) cmax = fmaxf(cmax, z[j]);
    const float new_m = fmaxf(run_m, cmax);  // finite: at least one real column in this chunk
    const float nm2 = new_m * LOG2E;
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    float t0 = 0.f, t1 = 0.f;
#pragma unroll
    for (int j = 0; j < 32; j += 4) {
      const float e0 = ex2_approx(fmaf(z[j + 0], LOG2E, -nm2));
      const float e1 = ex2_approx(fmaf(z[j + 1], LOG2E, -nm2));
      const float e2 = ex2_approx(fmaf(z[j + 2], LOG2E, -nm2));
      const float e3 = ex2_approx(fmaf(z[j + 3], LOG2E, -nm2));
      s0 += e0;
      s1 += e1;
      s2 += e2;
      s3 += e3;
      if constexpr (STATS) {  // masked columns hold -inf with e = 0: clamp so 0 * z stays 0
        t0 = fmaf(e0, fmaxf(z[j + 0], -3.0e38f), t0);
        t1 = fmaf(e1, fmaxf(z[j + 1], -3.0e38f), t1);
        t0 = fmaf(e2, fmaxf(z[j + 2], -3.0e38f), t0);
        t1 = fmaf(e3, fmaxf(z[j + 3], -3.0e38f), t1);
      }
    }
    const float rescale = ex2_approx((run_m - new_m) * LOG2E);
    run_s = run_s * rescale + ((s0 + s1) + (s2 + s3));
    if constexpr (STATS) run_ez = run_ez * rescale + (t0 + t1);
    run_m = new_m;
    if (lbl >= col0 && lbl < col0 + 32) {
      float t = 0.f;
#pragma unroll
      for (int j = 0; j < 32; ++j) t = (col0 + j == lbl) ? z[j] : t;
      if (row < M) p.tgt_logit[row] = t;
    }
  }

  __device__ __forceinline__ void tile_end(const TileCoord& tc, int) {
    if (row < M) {
      const size_t at = static_cast<size_t>(tc.tile_n) * p.row_pitch + row;
      p.partials[at] = make_float2(run_m, run_s);
      if constexpr (STATS) p.partials_ez[at] = run_ez;
    }
  }
};

// ---- bf16 output staging: registers -> swizzled smem -> TMA store ---------------------------------
// Each epilogue thread owns one output row and produces 32 consecutive columns per chunk.  Writing
// those straight to global memory makes every warp store touch 32 different lines; instead the warp
// stages a [32 rows x 64 columns] block in shared memory (128-byte rows, 128B swizzle: 16-byte piece
// j of row r sits at position j ^ (r & 7), conflict-free for the quarter-warp wavefronts) and one
// lane issues a TMA store for it.  The TMA unit clips the box at the tensor's edges, so ragged
// last tiles need no special handling.
struct Bf16TileStore {
  uint8_t* smem;  // this warp's EPI_WARP_SMEM_BYTES: two staging tiles
  int lane;
  int cur;          // staging tile being filled (0 / 1)
  uint64_t policy;  // L2 eviction priority of the stored lines (L2_EVICT_* in sm100_ptx.cuh)

  __device__ __forceinline__ void init(uint8_t* warp_smem, uint64_t l2_policy = L2_EVICT_NORMAL) {
    smem = warp_smem;
    lane = threadIdx.x & 31;
    cur = 0;
    policy = l2_policy ? l2_policy : L2_EVICT_NORMAL;
  }

  // `half` = which 32-column half of the 64-column block these values fill.
  __device__ __forceinline__ void stage(const float (&g)[32], int half) {
    if (half == 0) {  // this tile was last handed to the TMA unit two stores ago: only the latest store may still be reading
      if (lane == 0) tma_store_wait_read<1>();
      __syncwarp();
    }
    uint8_t* row = smem + cur * EPI_STAGE_TILE_BYTES + lane * 128;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      uint4 v;
      v.x = pack_bf16x2(g[q * 8 + 0], g[q * 8 + 1]);
      v.y = pack_bf16x2(g[q * 8 + 2], g[q * 8 + 3]);
      v.z = pack_bf16x2(g[q * 8 + 4], g[q * 8 + 5]);
      v.w = pack_bf16x2(g[q * 8 + 6], g[q * 8 + 7]);
      const int piece = (half * 4 + q) ^ (lane & 7);
      *reinterpret_cast<uint4*>(row + piece * 16) = v;
    }
  }

  // Store the staged block to global (col0 = first column of the 64-column block, row0 = first of
  // the warp's 32 rows).  Columns / rows beyond the tensor map's extents are not written.
  __device__ __forceinline__ void flush(const CUtensorMap* map, int col0, int row0) {
    fence_proxy_async_smem();
    __syncwarp();
    if (lane == 0) {
      tma_store_2d(map, smem + cur * EPI_STAGE_TILE_BYTES, col0, row0, policy);
      tma_store_commit();
    }
    cur ^= 1;
  }

  __device__ __forceinline__ void drain() {
    if (lane == 0) tma_store_wait<0>();
    __syncwarp();
  }
};

// ---- backward stage 1: dlogits tile -> bf16 scratch --------------------------------------------
struct DlogitsEpi {
  struct Params {
    const int32_t* labels;  // [M] (chunk-local)
    const float* lse;       // [M]
    const float* grad_out;  // device scalar
    const int64_t* denom;   // device scalar or nullptr
    float scale;
    int ignore_index;
    int round_bf16;
    CUtensorMap out_map;  // [rows_alloc][slab cols] bf16 view of this vocabulary slab of dlogits, box 64 x 32.
                          // Rows in [live rows, rows_alloc) are written as zeros because the dW GEMM
                          // contracts over them in whole 64-row blocks.
    int col_offset;       // first vocabulary column of the slab (labels are global column indices)
    uint64_t store_policy;  // L2 eviction priority of the dlogits stores (0 = normal)
  };
  const Params& p;
  const int M, N;
  float coef_all, coef, lse2;
  int lbl, row;  // lbl is slab-local: label - col_offset (never matches when outside the slab)
  Bf16TileStore st;

  __device__ DlogitsEpi(const Params& p_, const GemmShape& s, uint8_t* warp_smem) : p(p_), M(s.M), N(s.N) {
    float c = __ldg(p.grad_out) * p.scale;
    if (p.denom != nullptr) c = c / static_cast<float>(__ldg(p.denom));
    coef_all = c;
    st.init(warp_smem, p.store_policy);
  }
  __device__ __forceinline__ void finish() { st.drain(); }

  __device__ __forceinline__ void tile_begin(const TileCoord& tc, int row_in_tile) {
    row = tc.m0 + row_in_tile;
    int glbl = p.ignore_index;
    lse2 = 0.f;
    if (row < M) {
      glbl = __ldg(p.labels + row);
      lse2 = __ldg(p.lse + row) * LOG2E;
    }
    coef = (glbl != p.ignore_index) ? coef_all : 0.f;
    lbl = glbl - p.col_offset;
  }

  __device__ __forceinline__ void chunk(const TileCoord& tc, int row_in_tile, int c, const uint32_t (&acc)[32]) {
    const int col0 = tc.n0 + c;
    if (col0 >= N) return;  // warp-uniform
    float g[32];
    if (coef == 0.f) {  // ignored / dead row (or zero upstream gradient): exact zeros
#pragma unroll
      for (int j = 0; j < 32; ++j) g[j] = 0.f;
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        float z = __uint_as_float(acc[j]);
        if (p.round_bf16) z = round_to_bf16(z);
        float pr = ex2_approx(fmaf(z, LOG2E, -lse2));
        pr = (col0 + j == lbl) ? pr - 1.f : pr;
        g[j] = pr * coef;
      }
    }
    const int half = (c >> 5) & 1;
    st.stage(g, half);
    if (half == 1 || col0 + 32 >= N)  // block complete, or the slab ends inside it
      st.flush(&p.out_map, col0 - half * 32, tc.m0 + (row_in_tile & ~31));
  }

  __device__ __forceinline__ void tile_end(const TileCoord&, int) {}
};

// ---- plain store epilogue (dX, dW, exported GEMM) ----------------------------------------------
enum StoreMode { STORE_BF16 = 0, STORE_F32 = 1, ACCUM_F32 = 2, ACCUM_F32_OUT_BF16 = 3 };

struct StoreEpi {
  struct Params {
    CUtensorMap out_map;  // [rows][cols] bf16 output view, box 64 x 32 (modes 0, 3)
    float* acc;           // [M][ld] fp32 (modes 1, 2, 3); K-split ks uses acc + ks * split_stride
    size_t split_stride;  // elements between the per-split accumulators (0 when the GEMM is not split)
    int ld;
    int mode;
    float* tail_acc;      // [slots][tile rows][256] fp32: accumulators of the K pieces of tail tiles (GemmShape::tail_split)
    int tail_tile_rows;   // rows of one tile in that scratch (256 for the pair engine, 128 for the single-CTA one)
  };
  const Params& p;
  const int M, N;
  int row;
  Bf16TileStore st;

  __device__ StoreEpi(const Params& p_, const GemmShape& s, uint8_t* warp_smem) : p(p_), M(s.M), N(s.N) {
    st.init(warp_smem);
  }
  __device__ __forceinline__ void tile_begin(const TileCoord& tc, int row_in_tile) { row = tc.m0 + row_in_tile; }
  __device__ __forceinline__ void finish() { st.drain(); }

  __device__ __forceinline__ void chunk(const TileCoord& tc, int row_in_tile, int c, const uint32_t (&acc)[32]) {
    const int col0 = tc.n0 + c;
    if (tc.tail_slot >= 0) {  // a K piece of a tail tile: the raw accumulator goes to its scratch slot, whole tile
      float4* dst = reinterpret_cast<float4*>(
          p.tail_acc + (static_cast<size_t>(tc.tail_slot) * p.tail_tile_rows + tc.row_off + row_in_tile) * BLOCK_N + c);
#pragma unroll
      for (int q = 0; q < 8; ++q)
        dst[q] = make_float4(__uint_as_float(acc[q * 4 + 0]), __uint_as_float(acc[q * 4 + 1]),
                             __uint_as_float(acc[q * 4 + 2]), __uint_as_float(acc[q * 4 + 3]));
      return;
    }
    if (col0 >= N) return;  // warp-uniform
    float v[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = (row < M) ? __uint_as_float(acc[j]) : 0.f;
    if (p.mode != STORE_BF16 && row < M) {  // fp32 accumulator across launches (only used with long K: cheap)
      float* a = p.acc + static_cast<size_t>(tc.ks) * p.split_stride + static_cast<size_t>(row) * p.ld + col0;
      if (col0 + 32 <= N) {
        float4* a4 = reinterpret_cast<float4*>(a);
        if (p.mode != STORE_F32) {
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            float4 t = a4[q];
            v[q * 4 + 0] += t.x;
            v[q * 4 + 1] += t.y;
            v[q * 4 + 2] += t.z;
            v[q * 4 + 3] += t.w;
          }
        }
        if (p.mode != ACCUM_F32_OUT_BF16) {
#pragma unroll
          for (int q = 0; q < 8; ++q) a4[q] = make_float4(v[q * 4 + 0], v[q * 4 + 1], v[q * 4 + 2], v[q * 4 + 3]);
        }
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          if (col0 + j < N) {
            if (p.mode != STORE_F32) v[j] += a[j];
            if (p.mode != ACCUM_F32_OUT_BF16) a[j] = v[j];
          }
        }
      }
    }
    if (p.mode == STORE_BF16 || p.mode == ACCUM_F32_OUT_BF16) {
      const int half = (c >> 5) & 1;
      st.stage(v, half);
      if (half == 1 || col0 + 32 >= N) st.flush(&p.out_map, col0 - half * 32, tc.m0 + (row_in_tile & ~31));
    }
  }
  __device__ __forceinline__ void tile_end(const TileCoord&, int) {}
};

// =============================================================================================
// Row-wise kernels around the forward GEMM
// =============================================================================================

// Merge the per-vocabulary-tile (max, sumexp) partials of each row into lse, and emit the row's
// loss term (lse - target logit) or 0 for ignored rows.  With `partials_ez` also the row's softmax
// entropy and whether the target holds the row maximum (token accuracy).
constexpr int MERGE_ROWS = 32;   // rows per block
constexpr int MERGE_LANES = 8;   // threads per row, each folding every 8th vocabulary tile

__global__ void __launch_bounds__(MERGE_ROWS* MERGE_LANES)
    flce_merge_rows_kernel(const float2* __restrict__ partials, const float* __restrict__ partials_ez,
                           const float* __restrict__ tgt_logit, const int32_t* __restrict__ labels, int rows_cap,
                           const int* __restrict__ n_rows_dev, int tiles_n, int V, int ignore_index, float* __restrict__ lse,
                           float* __restrict__ row_loss, float* __restrict__ row_entropy,
                           unsigned char* __restrict__ row_correct) {
  const int M = n_rows_dev ? min(max(*n_rows_dev, 0), rows_cap) : rows_cap;
  const int r = threadIdx.x % MERGE_ROWS, g = threadIdx.x / MERGE_ROWS;
  const int row = blockIdx.x * MERGE_ROWS + r;
  // online merge of this lane's tiles: (m, s, ez) with s and ez relative to m
  float m = -INFINITY, sum = 0.f, ez = 0.f;
  if (row < M) {
    for (int t = g; t < tiles_n; t += MERGE_LANES) {
      const size_t at = static_cast<size_t>(t) * rows_cap + row;
      const float2 pr = partials[at];
      const float nm = fmaxf(m, pr.x);
      const float fo = exp2f((m - nm) * LOG2E), fn = exp2f((pr.x - nm) * LOG2E);
      sum = sum * fo + pr.y * fn;
      if (partials_ez) ez = ez * fo + partials_ez[at] * fn;
      m = nm;
    }
  }
  __shared__ float s_m[MERGE_LANES][MERGE_ROWS], s_s[MERGE_LANES][MERGE_ROWS], s_e[MERGE_LANES][MERGE_ROWS];
  s_m[g][r] = m;
  s_s[g][r] = sum;
  s_e[g][r] = ez;
  __syncthreads();
  if (g != 0 || row >= M) return;
  for (int k = 1; k < MERGE_LANES; ++k) {  // fixed order: deterministic
    const float om = s_m[k][r];
    if (om == -INFINITY) continue;  // that lane had no tiles
    const float nm = fmaxf(m, om);
    const float fo = exp2f((m - nm) * LOG2E), fn = exp2f((om - nm) * LOG2E);
    sum = sum * fo + s_s[k][r] * fn;
    ez = ez * fo + s_e[k][r] * fn;
    m = nm;
  }
  const float l = m + logf(sum);
  lse[row] = l;
  const int lbl = labels[row];
  const bool valid = lbl != ignore_index;
  const float tgt = tgt_logit[row];
  // A label outside [0,V) that is not ignore_index trips a device assert in the reference
  // (nll_loss); nothing here can raise, so the row -- and with it the loss -- turns NaN instead.
  const bool in_range = static_cast<unsigned int>(lbl) < static_cast<unsigned int>(V);
  row_loss[row] = valid ? (in_range ? (l - tgt) : __int_as_float(0x7fc00000)) : 0.f;
  if (partials_ez) {
    row_entropy[row] = valid ? (l - ez / sum) : 0.f;
    row_correct[row] = (valid && tgt >= m) ? 1 : 0;  // the target logit is the row maximum
  }
}

// One block: fixed-order (deterministic) sums of the row losses / entropies and counts.
__global__ void flce_reduce_loss_kernel(const float* __restrict__ row_loss, const float* __restrict__ row_entropy,
                                        const unsigned char* __restrict__ row_correct,
                                        const int32_t* __restrict__ labels, int rows_cap,
                                        const int* __restrict__ n_rows_dev, int ignore_index, float scale, int denom_mode,
                                        const int64_t* __restrict__ denom, float* __restrict__ loss,
                                        int64_t* __restrict__ n_valid, int64_t* __restrict__ n_correct,
                                        float* __restrict__ entropy_sum) {
  const int M = n_rows_dev ? min(max(*n_rows_dev, 0), rows_cap) : rows_cap;
  __shared__ double s_sum[1024];
  __shared__ double s_ent[1024];
  __shared__ int s_cnt[1024];
  __shared__ int s_cor[1024];
  double acc = 0.0, ent = 0.0;
  int cnt = 0, cor = 0;
  for (int i = threadIdx.x; i < M; i += blockDim.x) {
    acc += static_cast<double>(row_loss[i]);
    cnt += (labels[i] != ignore_index) ? 1 : 0;
    if (row_entropy) {
      ent += static_cast<double>(row_entropy[i]);
      cor += row_correct[i];
    }
  }
  s_sum[threadIdx.x] = acc;
  s_ent[threadIdx.x] = ent;
  s_cnt[threadIdx.x] = cnt;
  s_cor[threadIdx.x] = cor;
  __syncthreads();
  for (int off = blockDim.x >> 1; off > 0; off >>= 1) {
    if (threadIdx.x < off) {
      s_sum[threadIdx.x] += s_sum[threadIdx.x + off];
      s_ent[threadIdx.x] += s_ent[threadIdx.x + off];
      s_cnt[threadIdx.x] += s_cnt[threadIdx.x + off];
      s_cor[threadIdx.x] += s_cor[threadIdx.x + off];
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    float total = static_cast<float>(s_sum[0]) * scale;
    if (denom_mode == B200SFT_DENOM_DEVICE) total = total / static_cast<float>(*denom);
    if (denom_mode == B200SFT_DENOM_VALID) total = total / static_cast<float>(s_cnt[0]);  // 0/0 = NaN, as torch
    *loss = total;
    if (n_valid != nullptr) *n_valid = static_cast<int64_t>(s_cnt[0]);
    if (n_correct != nullptr) *n_correct = static_cast<int64_t>(s_cor[0]);
    if (entropy_sum != nullptr) *entropy_sum = static_cast<float>(s_ent[0]);
  }
}

// =============================================================================================
// Valid-row compaction: rows whose label is ignore_index add nothing to the loss, to dW, and get
// a zero dX row, so the three kernels below let every GEMM run on the live rows only.  The live
// count stays on the device (GemmShape::dyn_rows); nothing here synchronises the host.
// =============================================================================================

// The live-row list itself (row_index, labels_c, pos, n_live) comes from the label front-end kernel in
// elementwise.cu (launch_compact_labels: stable, one launch, an 8-CTA cluster for long batches).

// xc[j,:] = x[row_index[j],:] for j < n_live; rows up to the next multiple of 128 are zero-filled
// (GEMMs that contract over rows consume whole 64-row blocks).  One warp per row, 16-byte accesses.
__global__ void gather_rows_kernel(const __nv_bfloat16* __restrict__ x, const int32_t* __restrict__ row_index,
                                   const int* __restrict__ n_live, int N, int H, __nv_bfloat16* __restrict__ xc) {
  const int live = min(max(*n_live, 0), N);
  const int padded = min((live + 127) / 128 * 128, N);
  const int warps_per_block = blockDim.x >> 5, lane = threadIdx.x & 31;
  const int vec = H / 8;
  for (int j = blockIdx.x * warps_per_block + (threadIdx.x >> 5); j < padded; j += gridDim.x * warps_per_block) {
    uint4* dst = reinterpret_cast<uint4*>(xc + static_cast<size_t>(j) * H);
    if (j < live) {
      const uint4* src = reinterpret_cast<const uint4*>(x + static_cast<size_t>(row_index[j]) * H);
      for (int v = lane; v < vec; v += 32) dst[v] = src[v];
    } else {
      for (int v = lane; v < vec; v += 32) dst[v] = make_uint4(0, 0, 0, 0);
    }
  }
}

// dx[r,:] = dxc[pos[r],:] if pos[r] >= 0 else 0 -- every output row is written exactly once.
__global__ void scatter_rows_kernel(const __nv_bfloat16* __restrict__ dxc, const int32_t* __restrict__ pos, int N, int H,
                                    __nv_bfloat16* __restrict__ dx) {
  const int warps_per_block = blockDim.x >> 5, lane = threadIdx.x & 31;
  const int vec = H / 8;
  for (int r = blockIdx.x * warps_per_block + (threadIdx.x >> 5); r < N; r += gridDim.x * warps_per_block) {
    uint4* dst = reinterpret_cast<uint4*>(dx + static_cast<size_t>(r) * H);
    const int j = pos[r];
    if (j >= 0) {
      const uint4* src = reinterpret_cast<const uint4*>(dxc + static_cast<size_t>(j) * H);
      for (int v = lane; v < vec; v += 32) dst[v] = src[v];
    } else {
      for (int v = lane; v < vec; v += 32) dst[v] = make_uint4(0, 0, 0, 0);
    }
  }
}

// dX = bf16( sum over the K-split accumulators ).  The number of live splits is recomputed from the
// same inputs the GEMM kernel used (live rows -> tiles -> choose_k_splits), so nothing is read back.
__global__ void dx_combine_splits_kernel(const float* __restrict__ acc, size_t split_stride, int rows_cap, int H,
                                         const int* __restrict__ n_rows_dev, int tile_rows, int tiles_n, int workers,
                                         int max_splits, int min_splits, __nv_bfloat16* __restrict__ out) {
  const int rows = n_rows_dev ? min(max(*n_rows_dev, 0), rows_cap) : rows_cap;
  const int tiles = ((rows + tile_rows - 1) / tile_rows) * tiles_n;
  const int splits = choose_k_splits(tiles, workers, max_splits, min_splits);
  const size_t nvec = static_cast<size_t>(rows) * H / 4;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < nvec;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    float4 v = reinterpret_cast<const float4*>(acc)[i];
    for (int s = 1; s < splits; ++s) {
      const float4 t = reinterpret_cast<const float4*>(acc + s * split_stride)[i];
      v.x += t.x;
      v.y += t.y;
      v.z += t.z;
      v.w += t.w;
    }
    reinterpret_cast<uint2*>(out)[i] = make_uint2(pack_bf16x2(v.x, v.y), pack_bf16x2(v.z, v.w));
  }
}

// out[tile] = bf16( sum over the K pieces of a tail tile, in piece order ).  grid = (tail tiles, row groups of
// TAIL_COMBINE_ROWS); the tile's place follows from the same UnitPlan the GEMM used (tile_n fastest).  Rows / columns
// beyond the output are clipped.
constexpr int TAIL_COMBINE_ROWS = 8;
__global__ void __launch_bounds__(256)
    tail_combine_kernel(const float* __restrict__ tail_acc, UnitPlan plan, int tile_rows, int tiles_n, int M, int N, int ldc,
                        __nv_bfloat16* __restrict__ out) {
  const int t = blockIdx.x;  // tail tile index
  const int tile = plan.full_tiles + t;
  const int tile_m = tile / tiles_n, tile_n = tile - tile_m * tiles_n;
  const size_t tile_elems = static_cast<size_t>(tile_rows) * BLOCK_N;
  for (int i = threadIdx.x; i < TAIL_COMBINE_ROWS * (BLOCK_N / 4); i += blockDim.x) {
    const int r = blockIdx.y * TAIL_COMBINE_ROWS + i / (BLOCK_N / 4), c4 = (i % (BLOCK_N / 4)) * 4;
    if (r >= tile_rows) continue;
    const int row = tile_m * tile_rows + r, col = tile_n * BLOCK_N + c4;
    if (row >= M || col >= N) continue;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int k = 0; k < plan.tail_splits; ++k) {
      const float4 x = *reinterpret_cast<const float4*>(
          tail_acc + (static_cast<size_t>(t) * plan.tail_splits + k) * tile_elems + static_cast<size_t>(r) * BLOCK_N + c4);
      v.x += x.x;
      v.y += x.y;
      v.z += x.z;
      v.w += x.w;
    }
    __nv_bfloat16* o = out + static_cast<size_t>(row) * ldc + col;
    if (col + 4 <= N) {
      *reinterpret_cast<uint2*>(o) = make_uint2(pack_bf16x2(v.x, v.y), pack_bf16x2(v.z, v.w));
    } else {
      const float e[4] = {v.x, v.y, v.z, v.w};
      for (int j = 0; j < 4 && col + j < N; ++j) o[j] = __float2bfloat16_rn(e[j]);
    }
  }
}

// =============================================================================================
// Host side
// =============================================================================================
namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// bf16 2-D tensor map with 128-byte swizzle.  `inner` elements are contiguous; rows are `ld`
// elements apart.  Out-of-bounds box elements read as zero (they add nothing to the MMA).
int make_tmap(CUtensorMap* m, const void* base, uint64_t inner, uint64_t outer, uint64_t ld, uint32_t box_inner,
              uint32_t box_outer) {
  EncodeTiledFn fn = get_encode_fn();
  if (fn == nullptr) return fail(B200SFT_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
  if ((reinterpret_cast<uintptr_t>(base) & 15) != 0 || (ld * 2) % 16 != 0)
    return fail(B200SFT_ERR_INVALID, "TMA operand needs a 16-byte aligned base and row pitch (ld=%llu)",
                (unsigned long long)ld);
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {ld * 2};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(B200SFT_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return B200SFT_OK;
}

// Operand map: K-major = [rows][K] with K contiguous; MN-major = [K][rows] with rows contiguous.
int make_operand_tmap(CUtensorMap* m, const void* base, int rows, int K, int ld, bool mn_major, int block_rows) {
  if (!mn_major) return make_tmap(m, base, K, rows, ld, BLOCK_K, block_rows);
  return make_tmap(m, base, rows, K, ld, 64, BLOCK_K);
}

// Output view for the epilogue's TMA stores: [rows][cols] bf16, row pitch ld, box = 64 columns x 32 rows.
int make_store_tmap(CUtensorMap* m, const void* base, int rows, int cols, int ld) {
  return make_tmap(m, base, static_cast<uint64_t>(cols), static_cast<uint64_t>(rows), static_cast<uint64_t>(ld), 64, 32);
}

int current_device() {
  int dev = 0;
  cudaGetDevice(&dev);
  return (dev < 0 || dev >= 64) ? 0 : dev;
}

int num_sms() {
  static int sms[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (sms[dev] == 0) cudaDeviceGetAttribute(&sms[dev], cudaDevAttrMultiProcessorCount, dev);
  return sms[dev] > 0 ? sms[dev] : 148;
}

// Rasterisation group for the M-fastest (logit) GEMMs: 8192 rows of X (33.5 MB at
// H = 2048) stay L2-resident while every vocabulary tile is swept over them; W is re-streamed
// once per group.  Without it N = 65536 re-reads all of X from HBM for each of the 501 W panels.
constexpr int GROUP_M_ROWS = 8192;

// 1 = one CTA per 128x256 tile (cta_group::1); 2 = CTA pairs on 256x256 tiles (cta_group::2).
// Default 2: measured 5-6 % faster on the backward GEMMs and 4 % on the forward at N = 65536
// (profiles/r01_cta_pair_vs_single.log); results are bit-identical.
int g_cta_group = 2;

// BN = output tile width of the pair engine (256; 128 is implemented and tested but slower, see DX_TILE_N).
// Number of persistent workers (CTAs, or CTA pairs) a GEMM of this output shape is launched with.
int gemm_workers(int M, int N, int bn = BLOCK_N) {
  const bool paired = (g_cta_group == 2);
  const int tile_m = paired ? pair::TILE_M : BLOCK_M;
  const long long tiles = static_cast<long long>((M + tile_m - 1) / tile_m) * ((N + bn - 1) / bn);
  const int max_w = paired ? num_sms() / 2 : num_sms();
  return static_cast<int>(tiles < max_w ? tiles : max_w);
}

// Per-launch options of the engine beyond the problem shape (see GemmShape for what each field does).
struct GemmOpts {
  const int* dyn_rows = nullptr;
  int dyn_is_k = 0;
  int max_k_splits = 1, min_k_splits = 1;
  int ks_major = 0;
  uint64_t a_l2_policy = 0, b_l2_policy = 0;
  int k_pingpong = 0;
  int tail_split = 0;  // StoreEpi launches only: the caller supplies Params::tail_acc and runs launch_tail_combine after
};

template <class Epi, bool A_MN, bool B_MN, int BN = 256>
int launch_gemm(const void* a, int lda, const void* b, int ldb, int M, int N, int K, int m_fastest,
                const typename Epi::Params& ep, cudaStream_t stream, const GemmOpts& o = GemmOpts()) {
  if (M <= 0 || N <= 0 || K <= 0) return fail(B200SFT_ERR_INVALID, "empty GEMM %dx%dx%d", M, N, K);
  const bool paired = (g_cta_group == 2);
  const int bn = paired ? BN : BLOCK_N;  // the single-CTA engine has one tile shape
  CUtensorMap ta, tb;
  int rc = make_operand_tmap(&ta, a, M, K, lda, A_MN, BLOCK_M);
  if (rc != 0) return rc;
  rc = make_operand_tmap(&tb, b, N, K, ldb, B_MN, paired ? bn / 2 : BLOCK_N);
  if (rc != 0) return rc;
  const int tile_m = paired ? pair::TILE_M : BLOCK_M;
  GemmShape s;
  s.M = M;
  s.N = N;
  s.K = K;
  s.tiles_m = (M + tile_m - 1) / tile_m;
  s.tiles_n = (N + bn - 1) / bn;
  s.m_fastest = m_fastest;
  const int group = GROUP_M_ROWS / tile_m;
  s.group_m = s.tiles_m < group ? s.tiles_m : group;
  s.dyn_rows = o.dyn_rows;
  s.dyn_is_k = o.dyn_is_k;
  s.max_k_splits = o.max_k_splits;
  s.min_k_splits = o.min_k_splits;
  s.ks_major = o.ks_major;
  s.a_l2_policy = o.a_l2_policy;
  s.b_l2_policy = o.b_l2_policy;
  s.tail_split = o.tail_split;
  s.k_pingpong = o.k_pingpong;
  const long long tiles = static_cast<long long>(s.tiles_m) * s.tiles_n;
  if (!paired) {
    auto kern = gemm_tcgen05_kernel<Epi, A_MN, B_MN>;
    static bool smem_set[64] = {false};  // per template instantiation and device
    const int dev = current_device();
    if (!smem_set[dev]) {
      B200_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_SMEM_BYTES));
      smem_set[dev] = true;
    }
    const int grid = static_cast<int>(tiles < num_sms() ? tiles : num_sms());
    kern<<<grid, GEMM_THREADS, GEMM_SMEM_BYTES, stream>>>(ta, tb, s, ep);
    B200_CUDA_CHECK(cudaGetLastError());
    return B200SFT_OK;
  }
  auto kern = gemm_tcgen05_2cta_kernel<Epi, A_MN, B_MN, BN>;
  constexpr int smem_bytes = pair::Cfg<BN>::SMEM_BYTES;
  static bool smem_set[64] = {false};
  const int dev = current_device();
  if (!smem_set[dev]) {
    B200_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    smem_set[dev] = true;
  }
  const int max_pairs = num_sms() / 2;
  const int pairs = static_cast<int>(tiles < max_pairs ? tiles : max_pairs);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * pairs);
  cfg.blockDim = dim3(GEMM_THREADS);
  cfg.dynamicSmemBytes = smem_bytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  B200_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, ta, tb, s, ep));
  return B200SFT_OK;
}

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// Backward stages dlogits one vocabulary slab at a time: [N][slab_cols] bf16.  The default slab
// width keeps that scratch within ~1 GiB and splits V into equal slabs (multiples of 256 columns).
constexpr size_t DEFAULT_DLOGITS_SCRATCH_BYTES = size_t(1) << 30;
// dX = dlogits . W has few tiles with a very long K (32 x 8 tiles on 74 CTA pairs = 3.46 waves).  128-wide
// tiles would make it 6.92 waves, but measured 10.9 ms instead of 8.9 ms for the whole backward at
// N = 8192: operand traffic per flop rises 1.5x and the L2->SM path saturates.  So: 256.
constexpr int DX_TILE_N = 256;

int resolve_slab_cols(int N, int V, int chunk_cols) {
  const int v_up = (V + BLOCK_N - 1) / BLOCK_N * BLOCK_N;
  long long cmax;
  if (chunk_cols > 0) {
    cmax = (static_cast<long long>(chunk_cols) + BLOCK_N - 1) / BLOCK_N * BLOCK_N;
  } else {
    cmax = static_cast<long long>(DEFAULT_DLOGITS_SCRATCH_BYTES / (2ull * static_cast<size_t>(N))) / BLOCK_N * BLOCK_N;
    if (cmax < BLOCK_N) cmax = BLOCK_N;
  }
  if (cmax >= v_up) return V;  // one slab: no dX accumulation needed
  const int slabs = static_cast<int>((v_up + cmax - 1) / cmax);
  const int c = ((v_up + slabs - 1) / slabs + BLOCK_N - 1) / BLOCK_N * BLOCK_N;  // balanced
  return c;
}

int check_dims(int N, int H, int V) {
  if (N <= 0 || H <= 0 || V <= 0) return fail(B200SFT_ERR_INVALID, "N, H, V must be positive (got %d, %d, %d)", N, H, V);
  if (H % 8 != 0 || V % 8 != 0)
    return fail(B200SFT_ERR_INVALID, "H and V must be multiples of 8 for 16-byte TMA row pitches (got H=%d V=%d)", H, V);
  return B200SFT_OK;
}

}  // namespace
}  // namespace b200

using namespace b200;

namespace {

// Tuning switches, read from the environment at each call (measurement aid; the defaults are what ships):
//   B200SFT_L2_HINTS=mask   bit 0: evict_last on the operand every tile re-reads; bit 1: evict_first on the operand
//                           that is streamed once; bit 2: evict_first on the dlogits STORES (0 = no hints)
//   B200SFT_DX_KS_MAJOR=0   dX work units tile-major (the round-1 order) instead of K-split-major
//   B200SFT_K_PINGPONG=mask bit 0: dX, bit 1: dW -- alternate the K sweep direction from unit to unit (GemmShape)
//   B200SFT_DX_MIN_SPLITS=n force at least n K-splits of the dX GEMM (power of two); B200SFT_DX_MAX_SPLITS=n the cap (4)
int tune_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return (v != nullptr && *v != 0) ? atoi(v) : dflt;
}
uint64_t keep_policy() { return (tune_int("B200SFT_L2_HINTS", 1) & 1) ? L2_EVICT_LAST : 0; }
uint64_t stream_policy() { return (tune_int("B200SFT_L2_HINTS", 1) & 2) ? L2_EVICT_FIRST : 0; }

// Logit GEMMs (forward, dlogits recompute): X is swept by every vocabulary panel of a rasterisation group, so it is
// the operand to keep in the L2; W panels are read once per group.
GemmOpts logit_gemm_opts(const int32_t* n_rows_dev) {
  GemmOpts o;
  o.dyn_rows = n_rows_dev;
  o.a_l2_policy = keep_policy();
  return o;
}

struct FwdScratch {
  float2* partials;
  float* partials_ez;
  float* tgt;
  float* row_loss;
  float* row_entropy;
  unsigned char* row_correct;
};

size_t fwd_workspace_bytes(int N, int V, bool stats) {
  if (N <= 0 || V <= 0) return 0;
  const size_t tiles_n = (static_cast<size_t>(V) + BLOCK_N - 1) / BLOCK_N;
  size_t b = align_up(tiles_n * static_cast<size_t>(N) * sizeof(float2), 256);  // partials
  b += align_up(static_cast<size_t>(N) * sizeof(float), 256);                   // target logits
  b += align_up(static_cast<size_t>(N) * sizeof(float), 256);                   // row losses
  if (stats) {
    b += align_up(tiles_n * static_cast<size_t>(N) * sizeof(float), 256);  // sum exp*z partials
    b += align_up(static_cast<size_t>(N) * sizeof(float), 256);            // row entropies
    b += align_up(static_cast<size_t>(N), 256);                            // row correct flags
  }
  return b;
}

FwdScratch carve_fwd(void* workspace, int N, int V, bool stats) {
  const size_t tiles_n = (static_cast<size_t>(V) + BLOCK_N - 1) / BLOCK_N;
  uint8_t* ws = static_cast<uint8_t*>(workspace);
  FwdScratch f = {};
  f.partials = reinterpret_cast<float2*>(ws);
  ws += align_up(tiles_n * static_cast<size_t>(N) * sizeof(float2), 256);
  f.tgt = reinterpret_cast<float*>(ws);
  ws += align_up(static_cast<size_t>(N) * sizeof(float), 256);
  f.row_loss = reinterpret_cast<float*>(ws);
  ws += align_up(static_cast<size_t>(N) * sizeof(float), 256);
  if (stats) {
    f.partials_ez = reinterpret_cast<float*>(ws);
    ws += align_up(tiles_n * static_cast<size_t>(N) * sizeof(float), 256);
    f.row_entropy = reinterpret_cast<float*>(ws);
    ws += align_up(static_cast<size_t>(N) * sizeof(float), 256);
    f.row_correct = ws;
  }
  return f;
}

// Shared body of the forward entry points.  `n_rows_dev` (nullable) = live rows on the device.
int flce_fwd_impl(const void* x, const void* w, const int32_t* labels, int N, int H, int V, int ignore_index,
                  float scale, int denom_mode, const int64_t* denom, uint32_t flags, const int32_t* n_rows_dev,
                  float* loss, float* lse, int64_t* n_valid, bool stats, int64_t* n_correct, float* entropy_sum,
                  void* workspace, size_t workspace_bytes, cudaStream_t stream) {
  int rc = check_dims(N, H, V);
  if (rc != 0) return rc;
  if (!x || !w || !labels || !loss || !lse) return fail(B200SFT_ERR_INVALID, "null tensor pointer");
  if (stats && (!n_correct || !entropy_sum)) return fail(B200SFT_ERR_INVALID, "null statistics pointer");
  if (denom_mode == B200SFT_DENOM_DEVICE && denom == nullptr)
    return fail(B200SFT_ERR_INVALID, "denom_mode DEVICE needs a denom pointer");
  if (denom_mode < 0 || denom_mode > 2) return fail(B200SFT_ERR_INVALID, "bad denom_mode %d", denom_mode);
  const size_t need = fwd_workspace_bytes(N, V, stats);
  if (workspace == nullptr || workspace_bytes < need)
    return fail(B200SFT_ERR_WORKSPACE, "flce_fwd workspace too small: have %zu, need %zu", workspace_bytes, need);
  const FwdScratch f = carve_fwd(workspace, N, V, stats);
  const int tiles_n = (V + BLOCK_N - 1) / BLOCK_N;

  // Rows whose label is ignored (or outside [0,V): those make the loss NaN) never get a target logit written.
  B200_CUDA_CHECK(cudaMemsetAsync(f.tgt, 0, static_cast<size_t>(N) * sizeof(float), stream));

  if (stats) {
    FwdSoftmaxEpi<true>::Params ep = {labels, f.partials, f.partials_ez, f.tgt, N,
                                      (flags & B200SFT_FLAG_ROUND_LOGITS_BF16) ? 1 : 0};
    rc = launch_gemm<FwdSoftmaxEpi<true>, false, false>(x, H, w, H, N, V, H, /*m_fastest=*/1, ep, stream,
                                                        logit_gemm_opts(n_rows_dev));
  } else {
    FwdSoftmaxEpi<false>::Params ep = {labels, f.partials, nullptr, f.tgt, N,
                                       (flags & B200SFT_FLAG_ROUND_LOGITS_BF16) ? 1 : 0};
    rc = launch_gemm<FwdSoftmaxEpi<false>, false, false>(x, H, w, H, N, V, H, /*m_fastest=*/1, ep, stream,
                                                         logit_gemm_opts(n_rows_dev));
  }
  if (rc != 0) return rc;

  flce_merge_rows_kernel<<<(N + MERGE_ROWS - 1) / MERGE_ROWS, MERGE_ROWS * MERGE_LANES, 0, stream>>>(f.partials, f.partials_ez, f.tgt, labels, N, n_rows_dev,
                                                               tiles_n, V, ignore_index, lse, f.row_loss, f.row_entropy,
                                                               f.row_correct);
  B200_CUDA_CHECK(cudaGetLastError());
  flce_reduce_loss_kernel<<<1, 1024, 0, stream>>>(f.row_loss, f.row_entropy, f.row_correct, labels, N, n_rows_dev,
                                                  ignore_index, scale, denom_mode, denom, loss, n_valid, n_correct,
                                                  entropy_sum);
  B200_CUDA_CHECK(cudaGetLastError());
  return B200SFT_OK;
}

// K-splits reserved for the dX GEMM: it has (N/tile rows) x (H/256) tiles, each with a very long K, so
// when that count is small relative to the workers, wave quantisation costs up to 14 %.  Up to 4
// splits (each with its own fp32 accumulator) when the GEMM has fewer than 16 tiles per worker.
int dx_max_k_splits(int N, int H) {
  const int workers = gemm_workers(N, H, DX_TILE_N);
  const int tile_m = (g_cta_group == 2) ? pair::TILE_M : BLOCK_M;
  const long long tiles = static_cast<long long>((N + tile_m - 1) / tile_m) * ((H + DX_TILE_N - 1) / DX_TILE_N);
  return tiles >= 16LL * workers ? 1 : tune_int("B200SFT_DX_MAX_SPLITS", 4);
}

// Lower bound on the dX K-splits (tuning switch only).  Measured on B200 at N = 8192 (profiles/r02_bwd_tuning.md):
// split-major order with the 2 splits the fill rule picks cuts the dX GEMM's DRAM reads from 3.1-3.2 GB to
// 2.2-2.3 GB per 64 Ki-column slab and its time from 1.46 to 1.41 ms, exactly the traffic of "the tiles in flight
// share their operands, nothing survives from one wave to the next"; forcing 4 or 8 splits to make the W range
// small enough to survive between waves did not lower the reads further and costs an fp32 [N,H] buffer per split.
int dx_min_k_splits(int N, int H, int slab_cols) {
  (void)slab_cols;
  const int maxs = dx_max_k_splits(N, H);
  const int forced = tune_int("B200SFT_DX_MIN_SPLITS", 0);
  return forced > 0 ? (forced < maxs ? forced : maxs) : 1;
}

// dW GEMM of one slab: [cols, H] output, tile_n fastest.  Its last wave is usually a sliver (2008 tiles on 74 pairs),
// so the tiles of that wave are K-split into one short wave (GemmShape::tail_split); B200SFT_DW_TAIL_SPLIT=0 turns it off.
bool dw_tail_split_on() { return tune_int("B200SFT_DW_TAIL_SPLIT", 1) != 0; }
UnitPlan dw_unit_plan(int cols, int H) {
  const int tile_m = (g_cta_group == 2) ? pair::TILE_M : BLOCK_M;
  const int tiles = ((cols + tile_m - 1) / tile_m) * ((H + BLOCK_N - 1) / BLOCK_N);
  return make_unit_plan(tiles, gemm_workers(cols, H), /*num_kb=*/0, 1, 1, 0, dw_tail_split_on() ? 1 : 0);
}
// One fp32 tile per worker at most (tail tiles x pieces <= workers): 148 x 128 x 256 x 4 B = 19.4 MB.
size_t dw_tail_scratch_bytes() { return align_up(static_cast<size_t>(num_sms()) * BLOCK_M * BLOCK_N * sizeof(float), 1024); }

}  // namespace

// =============================================================================================
// C ABI
// =============================================================================================
extern "C" {

size_t b200sft_flce_fwd_workspace_bytes(int N, int H, int V) {
  (void)H;
  return fwd_workspace_bytes(N, V, false);
}

size_t b200sft_flce_fwd_stats_workspace_bytes(int N, int H, int V) {
  (void)H;
  return fwd_workspace_bytes(N, V, true);
}

int b200sft_flce_fwd(const void* x, const void* w, const int32_t* labels, int N, int H, int V, int ignore_index,
                     float scale, int denom_mode, const int64_t* denom, uint32_t flags, const int32_t* n_rows_dev,
                     float* loss, float* lse, int64_t* n_valid, void* workspace, size_t workspace_bytes, void* stream) {
  return flce_fwd_impl(x, w, labels, N, H, V, ignore_index, scale, denom_mode, denom, flags, n_rows_dev, loss, lse,
                       n_valid, false, nullptr, nullptr, workspace, workspace_bytes, static_cast<cudaStream_t>(stream));
}

int b200sft_flce_fwd_stats(const void* x, const void* w, const int32_t* labels, int N, int H, int V, int ignore_index,
                           float scale, int denom_mode, const int64_t* denom, uint32_t flags,
                           const int32_t* n_rows_dev, float* loss, float* lse, int64_t* n_valid, int64_t* n_correct,
                           float* entropy_sum, void* workspace, size_t workspace_bytes, void* stream) {
  return flce_fwd_impl(x, w, labels, N, H, V, ignore_index, scale, denom_mode, denom, flags, n_rows_dev, loss, lse,
                       n_valid, true, n_correct, entropy_sum, workspace, workspace_bytes,
                       static_cast<cudaStream_t>(stream));
}

int b200sft_compact_rows(const void* x, const int32_t* labels, int N, int H, int ignore_index, void* xc,
                         int32_t* labels_c, int32_t* row_index, int32_t* pos, int32_t* n_live, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (N <= 0 || H <= 0 || H % 8 != 0) return fail(B200SFT_ERR_INVALID, "compact_rows: bad shape N=%d H=%d", N, H);
  if (!x || !labels || !xc || !labels_c || !row_index || !pos || !n_live)
    return fail(B200SFT_ERR_INVALID, "null tensor pointer");
  const int rc = launch_compact_labels(labels, N, ignore_index, row_index, labels_c, pos, n_live, stream);
  if (rc != 0) return rc;
  const int warps = 8;
  int grid = (N + warps - 1) / warps;
  if (grid > num_sms() * 8) grid = num_sms() * 8;
  gather_rows_kernel<<<grid, warps * 32, 0, stream>>>(static_cast<const __nv_bfloat16*>(x), row_index, n_live, N, H,
                                                       static_cast<__nv_bfloat16*>(xc));
  B200_CUDA_CHECK(cudaGetLastError());
  return B200SFT_OK;
}

int b200sft_gather_rows(const void* x, const int32_t* row_index, const int32_t* n_live, int N, int H, void* xc,
                        void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (N <= 0 || H <= 0 || H % 8 != 0) return fail(B200SFT_ERR_INVALID, "gather_rows: bad shape N=%d H=%d", N, H);
  if (!x || !row_index || !n_live || !xc) return fail(B200SFT_ERR_INVALID, "null tensor pointer");
  const int warps = 8;
  int grid = (N + warps - 1) / warps;
  if (grid > num_sms() * 8) grid = num_sms() * 8;
  gather_rows_kernel<<<grid, warps * 32, 0, stream>>>(static_cast<const __nv_bfloat16*>(x), row_index, n_live, N, H,
                                                       static_cast<__nv_bfloat16*>(xc));
  B200_CUDA_CHECK(cudaGetLastError());
  return B200SFT_OK;
}

int b200sft_scatter_rows(const void* dxc, const int32_t* pos, int N, int H, void* dx, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (N <= 0 || H <= 0 || H % 8 != 0) return fail(B200SFT_ERR_INVALID, "scatter_rows: bad shape N=%d H=%d", N, H);
  if (!dxc || !pos || !dx) return fail(B200SFT_ERR_INVALID, "null tensor pointer");
  const int warps = 8;
  int grid = (N + warps - 1) / warps;
  if (grid > num_sms() * 8) grid = num_sms() * 8;
  scatter_rows_kernel<<<grid, warps * 32, 0, stream>>>(static_cast<const __nv_bfloat16*>(dxc), pos, N, H,
                                                        static_cast<__nv_bfloat16*>(dx));
  B200_CUDA_CHECK(cudaGetLastError());
  return B200SFT_OK;
}

size_t b200sft_flce_bwd_workspace_bytes(int N, int H, int V, int chunk_cols) {
  if (N <= 0 || H <= 0 || V <= 0) return 0;
  const int c = resolve_slab_cols(N, V, chunk_cols);
  size_t b = align_up(static_cast<size_t>(N) * c * sizeof(__nv_bfloat16), 1024);  // dlogits slab
  const int splits = dx_max_k_splits(N, H);
  if (c < V || splits > 1)  // fp32 dX accumulator(s): across slabs and / or across K-splits
    b += static_cast<size_t>(splits) * align_up(static_cast<size_t>(N) * H * sizeof(float), 1024);
  b += dw_tail_scratch_bytes();  // K pieces of the dW GEMM's tail tiles
  return b;
}

int b200sft_flce_bwd_launch_count(int N, int H, int V, int chunk_cols, int need_dx, int need_dw) {
  if (N <= 0 || H <= 0 || V <= 0) return 0;
  const int slab = resolve_slab_cols(N, V, chunk_cols);
  const int num_slabs = (V + slab - 1) / slab;
  const int combine = (need_dx && dx_max_k_splits(N, H) > 1) ? 1 : 0;
  int tails = 0;
  for (int si = 0; need_dw && si < num_slabs; ++si) {
    const int cols = (V - si * slab) < slab ? (V - si * slab) : slab;
    tails += dw_unit_plan(cols, H).tail_splits > 1 ? 1 : 0;
  }
  return num_slabs * (1 + (need_dx ? 1 : 0) + (need_dw ? 1 : 0)) + combine + tails;
}

int b200sft_flce_bwd(const void* x, const void* w, const int32_t* labels, const float* lse, const float* grad_out,
                     int N, int H, int V, int ignore_index, float scale, int denom_mode, const int64_t* denom,
                     uint32_t flags, int chunk_cols, const int32_t* n_rows_dev, void* dx, void* dw, void* workspace,
                     size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  int rc = check_dims(N, H, V);
  if (rc != 0) return rc;
  if (!x || !w || !labels || !lse || !grad_out) return fail(B200SFT_ERR_INVALID, "null tensor pointer");
  if (!dx && !dw) return fail(B200SFT_ERR_INVALID, "at least one of dx / dw must be requested");
  if (denom_mode == B200SFT_DENOM_VALID)
    return fail(B200SFT_ERR_INVALID,
                "backward takes the valid-row count as a device scalar: pass n_valid from the forward call with "
                "denom_mode DEVICE");
  if (denom_mode == B200SFT_DENOM_DEVICE && denom == nullptr)
    return fail(B200SFT_ERR_INVALID, "denom_mode DEVICE needs a denom pointer");
  const size_t need = b200sft_flce_bwd_workspace_bytes(N, H, V, chunk_cols);
  if (workspace == nullptr || workspace_bytes < need)
    return fail(B200SFT_ERR_WORKSPACE, "flce_bwd workspace too small: have %zu, need %zu", workspace_bytes, need);

  const int slab = resolve_slab_cols(N, V, chunk_cols);
  const int num_slabs = (V + slab - 1) / slab;
  const int dx_splits = dx_max_k_splits(N, H);
  const int dx_min_splits = dx_min_k_splits(N, H, slab);
  const size_t dx_split_stride = align_up(static_cast<size_t>(N) * H * sizeof(float), 1024) / sizeof(float);
  uint8_t* ws = static_cast<uint8_t*>(workspace);
  __nv_bfloat16* dlog = reinterpret_cast<__nv_bfloat16*>(ws);
  ws += align_up(static_cast<size_t>(N) * slab * sizeof(__nv_bfloat16), 1024);
  float* dx_acc = (num_slabs > 1 || dx_splits > 1) ? reinterpret_cast<float*>(ws) : nullptr;
  if (dx_acc != nullptr) ws += static_cast<size_t>(dx_splits) * align_up(static_cast<size_t>(N) * H * sizeof(float), 1024);
  float* dw_tail_acc = reinterpret_cast<float*>(ws);

  const __nv_bfloat16* wb = static_cast<const __nv_bfloat16*>(w);
  __nv_bfloat16* dwb = static_cast<__nv_bfloat16*>(dw);

  for (int si = 0; si < num_slabs; ++si) {
    const int v0 = si * slab;
    const int cols = (V - v0) < slab ? (V - v0) : slab;
    const __nv_bfloat16* ws_rows = wb + static_cast<size_t>(v0) * H;  // W[v0 : v0+cols, :]

    // (1) dlogits[N,cols] = (softmax(X.W^T)[:, v0:v0+cols] - onehot) * coef   (recompute, epilogue-fused)
    DlogitsEpi::Params dp;
    dp.labels = labels;
    dp.lse = lse;
    dp.grad_out = grad_out;
    dp.denom = (denom_mode == B200SFT_DENOM_DEVICE) ? denom : nullptr;
    dp.scale = scale;
    dp.ignore_index = ignore_index;
    dp.round_bf16 = (flags & B200SFT_FLAG_ROUND_LOGITS_BF16) ? 1 : 0;
    rc = make_store_tmap(&dp.out_map, dlog, N, cols, slab);
    if (rc != 0) return rc;
    dp.col_offset = v0;
    dp.store_policy = (tune_int("B200SFT_L2_HINTS", 1) & 4) ? L2_EVICT_FIRST : 0;
    rc = launch_gemm<DlogitsEpi, false, false>(x, H, ws_rows, H, N, cols, H, /*m_fastest=*/1, dp, stream,
                                               logit_gemm_opts(n_rows_dev));
    if (rc != 0) return rc;

    // (2) dX[N,H] (+)= dlogits[N,cols] . W[v0:v0+cols, :]     A K-major, B = W rows read MN-major
    if (dx != nullptr) {
    StoreEpi::Params sx;
    rc = make_store_tmap(&sx.out_map, dx, N, H, H);
    if (rc != 0) return rc;
    sx.tail_acc = nullptr;
    sx.tail_tile_rows = 0;
    sx.acc = dx_acc;
    sx.split_stride = dx_splits > 1 ? dx_split_stride : 0;
    sx.ld = H;
    if (dx_splits > 1)  // every K-split keeps its own fp32 chain across the slabs; summed after the loop
      sx.mode = (si == 0) ? STORE_F32 : ACCUM_F32;
    else if (num_slabs == 1)
      sx.mode = STORE_BF16;
    else if (si == 0)
      sx.mode = STORE_F32;
    else if (si == num_slabs - 1)
      sx.mode = ACCUM_F32_OUT_BF16;
    else
      sx.mode = ACCUM_F32;
    GemmOpts ox;  // dlogits is streamed once per K range; the W rows of that range are what every tile re-reads
    ox.dyn_rows = n_rows_dev;
    ox.max_k_splits = dx_splits;
    ox.min_k_splits = dx_min_splits;
    ox.ks_major = tune_int("B200SFT_DX_KS_MAJOR", 1);
    ox.k_pingpong = tune_int("B200SFT_K_PINGPONG", 3) & 1;
    ox.a_l2_policy = stream_policy();
    ox.b_l2_policy = keep_policy();
    rc = launch_gemm<StoreEpi, false, true, DX_TILE_N>(dlog, slab, ws_rows, H, N, H, cols, /*m_fastest=*/0, sx, stream, ox);
    if (rc != 0) return rc;
    }

    // (3) dW[v0:v0+cols, :] = dlogits^T[cols,N] . X[N,H]      both operands read MN-major; K spans every
    //     row, so each dW slab is final after one GEMM (no accumulation across slabs)
    if (dw == nullptr) continue;  // frozen head: only dX was asked for
    StoreEpi::Params sw;
    rc = make_store_tmap(&sw.out_map, dwb + static_cast<size_t>(v0) * H, cols, H, H);
    if (rc != 0) return rc;
    sw.acc = nullptr;
    sw.split_stride = 0;
    sw.ld = H;
    sw.mode = STORE_BF16;
    sw.tail_acc = dw_tail_acc;
    sw.tail_tile_rows = (g_cta_group == 2) ? pair::TILE_M : BLOCK_M;
    GemmOpts ow;  // every wave of vocabulary tiles contracts over all of X: keep X, let dlogits stream past it
    ow.dyn_rows = n_rows_dev;
    ow.dyn_is_k = 1;
    ow.a_l2_policy = stream_policy();
    ow.b_l2_policy = keep_policy();
    ow.tail_split = dw_tail_split_on() ? 1 : 0;
    ow.k_pingpong = (tune_int("B200SFT_K_PINGPONG", 3) >> 1) & 1;
    rc = launch_gemm<StoreEpi, true, true>(dlog, slab, x, H, cols, H, N, /*m_fastest=*/0, sw, stream, ow);
    if (rc != 0) return rc;
    const UnitPlan wp = dw_unit_plan(cols, H);
    if (wp.tail_splits > 1) {  // add up the K pieces of the last wave's tiles
      tail_combine_kernel<<<dim3(wp.num_tiles - wp.full_tiles, sw.tail_tile_rows / TAIL_COMBINE_ROWS), 256, 0, stream>>>(
          dw_tail_acc, wp, sw.tail_tile_rows, (H + BLOCK_N - 1) / BLOCK_N, cols, H, H, dwb + static_cast<size_t>(v0) * H);
      B200_CUDA_CHECK(cudaGetLastError());
    }
  }
  if (dx_splits > 1 && dx != nullptr) {
    const int tile_m = (g_cta_group == 2) ? pair::TILE_M : BLOCK_M;
    const size_t nvec = static_cast<size_t>(N) * H / 4;
    size_t blocks = (nvec + 255) / 256;
    if (blocks > static_cast<size_t>(num_sms()) * 8) blocks = static_cast<size_t>(num_sms()) * 8;
    dx_combine_splits_kernel<<<static_cast<unsigned int>(blocks), 256, 0, stream>>>(
        dx_acc, dx_split_stride, N, H, n_rows_dev, tile_m, (H + DX_TILE_N - 1) / DX_TILE_N,
        gemm_workers(N, H, DX_TILE_N), dx_splits, dx_min_splits, static_cast<__nv_bfloat16*>(dx));
    B200_CUDA_CHECK(cudaGetLastError());
  }
  return B200SFT_OK;
}

int b200sft_set_cta_group(int cta_group) {
  if (cta_group != 1 && cta_group != 2) return fail(B200SFT_ERR_INVALID, "cta_group must be 1 or 2 (got %d)", cta_group);
  g_cta_group = cta_group;
  return B200SFT_OK;
}

int b200sft_gemm_bf16(const void* a, const void* b, void* c, int M, int N, int K, int lda, int ldb, int ldc,
                      int a_mn_major, int b_mn_major, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (!a || !b || !c) return fail(B200SFT_ERR_INVALID, "null tensor pointer");
  if (ldc % 8 != 0) return fail(B200SFT_ERR_INVALID, "ldc must be a multiple of 8");
  StoreEpi::Params sp;
  int rc = make_store_tmap(&sp.out_map, c, M, N, ldc);
  if (rc != 0) return rc;
  sp.acc = nullptr;
  sp.split_stride = 0;
  sp.ld = ldc;
  sp.mode = STORE_BF16;
  sp.tail_acc = nullptr;
  sp.tail_tile_rows = 0;
  if (!a_mn_major && !b_mn_major) return launch_gemm<StoreEpi, false, false>(a, lda, b, ldb, M, N, K, 0, sp, stream);
  if (!a_mn_major && b_mn_major) return launch_gemm<StoreEpi, false, true>(a, lda, b, ldb, M, N, K, 0, sp, stream);
  if (a_mn_major && !b_mn_major) return launch_gemm<StoreEpi, true, false>(a, lda, b, ldb, M, N, K, 0, sp, stream);
  return launch_gemm<StoreEpi, true, true>(a, lda, b, ldb, M, N, K, 0, sp, stream);
}

}  // extern "C"

// =============================================================================================
// Forward entry point that allocates nothing
// =============================================================================================
// Scratch for b200sft_flce_fwd_noalloc lives in the library image itself: every GPU that loads
// libb200sft.so reserves B200SFT_NOALLOC_SCRATCH_BYTES of HBM for it until the process exits.
// `g_noalloc_scratch` below is only a host-side handle for that array -- the device address is
// looked up with cudaGetSymbolAddress, once per device, before the first launch.
static __device__ __align__(256) unsigned char g_noalloc_scratch[B200SFT_NOALLOC_SCRATCH_BYTES];

namespace {

constexpr int kMaxDevices = 64;
void* g_noalloc_ptr[kMaxDevices] = {nullptr};

// Device address of the scratch array on the current device (resolved on first use per device).
int noalloc_scratch_ptr(void** out) {
  int dev = -1;
  B200_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return fail(B200SFT_ERR_UNSUPPORTED, "device ordinal %d out of range", dev);
  void* p = __atomic_load_n(&g_noalloc_ptr[dev], __ATOMIC_ACQUIRE);
  if (p == nullptr) {
    cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess || p == nullptr)
      return fail(B200SFT_ERR_CUDA, "cudaGetSymbolAddress(noalloc scratch) failed on device %d: %s", dev,
                  cudaGetErrorString(e));
    __atomic_store_n(&g_noalloc_ptr[dev], p, __ATOMIC_RELEASE);  // racing first calls resolve the same address
  }
  *out = p;
  return B200SFT_OK;
}

}  // namespace

extern "C" int b200sft_flce_fwd_noalloc(const void* x, const void* w, const int32_t* labels, int N, int H, int V,
                                        int ignore_index, float scale, int denom_mode, const int64_t* denom,
                                        uint32_t flags, float* loss, float* lse, int64_t* n_valid, void* stream) {
  const size_t need = b200sft_flce_fwd_workspace_bytes(N, H, V);
  if (need > static_cast<size_t>(B200SFT_NOALLOC_SCRATCH_BYTES))
    return fail(B200SFT_ERR_WORKSPACE,
                "b200sft_flce_fwd_noalloc: N=%d x V=%d needs %zu bytes of scratch but the built-in scratch holds %zu; "
                "split the rows into smaller calls or use b200sft_flce_fwd with caller-provided scratch",
                N, V, need, static_cast<size_t>(B200SFT_NOALLOC_SCRATCH_BYTES));
  void* scratch = nullptr;
  int rc = noalloc_scratch_ptr(&scratch);
  if (rc != 0) return rc;
  // Same kernels as b200sft_flce_fwd.  The array keeps whatever the previous call left in it: the
  // per-tile partials and row losses are fully rewritten for this call's N and V, and the
  // target-logit slots (the one region that must start at zero) are cleared in-stream by the
  // forward path itself before the GEMM.
  return b200sft_flce_fwd(x, w, labels, N, H, V, ignore_index, scale, denom_mode, denom, flags, /*n_rows_dev=*/nullptr,
                          loss, lse, n_valid, scratch, static_cast<size_t>(B200SFT_NOALLOC_SCRATCH_BYTES), stream);
}
