// Persistent, warp-specialised tcgen05 GEMM engine for sm_100a.
//
//   D[M,N] (fp32, TMEM) = A[M,K] (bf16) * B[N,K]^T (bf16)
//
// Two variants share the structure, the work-unit schedule and the epilogue functors:
//   gemm_tcgen05_kernel       one CTA per SM on 128x256 tiles (cta_group::1), 4-deep ring;
//   gemm_tcgen05_2cta_kernel  clusters of two CTAs on 256x256 tiles (cta_group::2), 6-deep ring -- the
//                             default (see the second half of this file).
// Warp 0 streams A/B k-blocks into the shared-memory ring with TMA (128-byte swizzle), warp 1 issues
// tcgen05.mma from one thread into one of two 256-column TMEM accumulator stages, warps 2..5 drain
// the other stage with tcgen05.ld and run the fused epilogue, so the epilogue of work unit i overlaps
// the MMAs of unit i+1.  A work unit is an output tile, or one K-split of it (choose_k_splits).  The
// epilogue is a functor; the LM-head/cross-entropy kernels differ only there (flce.cu).
//
// Either operand may be K-major (K contiguous in global memory) or MN-major (M or N contiguous):
// the data-gradient GEMMs read the same row-major tensors transposed, and MN-major descriptors
// avoid materialising transposes in HBM.
#pragma once
#include "sm100_ptx.cuh"

namespace b200 {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_N = 256;
constexpr int BLOCK_K = 64;  // 64 bf16 = one 128-byte swizzle row
constexpr int UMMA_K = 16;
constexpr int NUM_STAGES = 4;
constexpr int NUM_ACC_STAGES = 2;
constexpr int A_STAGE_BYTES = BLOCK_M * BLOCK_K * 2;  // 16 KiB
constexpr int B_STAGE_BYTES = BLOCK_N * BLOCK_K * 2;  // 32 KiB
constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
constexpr int TMEM_COLS = NUM_ACC_STAGES * BLOCK_N;  // 512: the whole TMEM of the SM
constexpr int NUM_EPI_WARPS = 4;
constexpr int GEMM_THREADS = 32 * (2 + NUM_EPI_WARPS);
constexpr int SMEM_BARRIER_BYTES = 256;
// Per-epilogue-warp staging for TMA stores: two tiles of 32 rows x 128 B (64 bf16 columns), 128B-swizzled, used
// alternately so that the warp fills one while the TMA unit still reads the other.
constexpr int EPI_STAGE_TILE_BYTES = 32 * 128;
constexpr int EPI_WARP_SMEM_BYTES = 2 * EPI_STAGE_TILE_BYTES;
constexpr int EPI_SMEM_BYTES = NUM_EPI_WARPS * EPI_WARP_SMEM_BYTES;
constexpr int GEMM_SMEM_BYTES =
    NUM_STAGES * STAGE_BYTES + EPI_SMEM_BYTES + 1024 /*alignment slack*/ + SMEM_BARRIER_BYTES;

struct GemmShape {
  int M, N, K;
  int tiles_m, tiles_n;
  int m_fastest;  // tile order: 1 = consecutive CTAs walk M for a fixed N tile (shares the B panel in L2)
  int group_m;    // m_fastest only: M tiles per rasterisation group.  All N tiles of a group are finished
                  // before the next group starts, so the group's A rows stay resident in L2.
  // Optional live row count on the device (valid-row compaction: the host never learns it).  When
  // non-null it replaces M (dyn_is_k = 0) or K (dyn_is_k = 1); the static value is the capacity the
  // tensor maps were built for.
  const int* dyn_rows;
  int dyn_is_k;
  // Split-K: when a GEMM has too few tiles to fill the workers evenly (dX: 256 tiles of K = 64k on 74 CTA
  // pairs = 3.46 waves), each tile's K range is cut into `choose_k_splits(...)` <= max_k_splits pieces that
  // become separate work units.  Every piece accumulates into its own fp32 buffer (TileCoord::ks selects
  // it in the epilogue), so the result stays deterministic; the caller sums the buffers afterwards.
  int max_k_splits;
  int min_k_splits;
  // Work-unit order when the GEMM is K-split: 0 = unit -> (tile, split) with the split fastest; 1 = split-major,
  // unit -> (split, tile): every tile of the GEMM is visited for one K range before the next range starts, so
  // the tiles in flight at any time all read the SAME K range of both operands (dX: the same rows of W[slab],
  // which then stay L2-resident across waves instead of being re-streamed for every wave of M tiles).
  int ks_major;
  // L2 eviction priority of the TMA loads of each operand (L2_EVICT_* in sm100_ptx.cuh; 0 = normal).
  uint64_t a_l2_policy, b_l2_policy;
  // Tail splitting (unsplit, tile_n-fastest GEMMs with a static tile count, i.e. dW): when the last wave of tiles
  // would leave most workers idle (2008 tiles on 74 CTA pairs: a 28th wave that is 13 % full), each of those last
  // tiles is cut into workers / tail K pieces that run as one short wave; every piece leaves its fp32 accumulator in
  // a scratch slot (TileCoord::tail_slot) and a small kernel adds the slots in fixed order.  See make_unit_plan.
  int tail_split;
  // Alternate the direction of the K sweep from one work unit of a worker to its next (all workers of a wave hold the
  // same parity): the K range the previous wave touched last is the one the next wave touches first, so whatever part
  // of the shared operand the L2 still holds is re-used instead of being evicted just before it is needed again.
  int k_pingpong;
};

struct TileCoord {
  int m0, n0, tile_m, tile_n;
  int ks;         // K-split index of this work unit (0 when the GEMM is not split)
  int tail_slot;  // >= 0: this unit is a K piece of a tail tile; its accumulator goes to that scratch slot
  int row_off;    // first row of this CTA inside the tile (cta_group::2: 0 or 128)
};

// Smallest power of two in [min_splits, max_splits] for which `tiles * splits` work units fill `workers` to
// >= 95 % in whole rounds.  A pure function of its arguments: kernels, the host and the combine kernel agree.
// min_splits > 1 is used with split-major order to keep one K range of the re-read operand L2-sized.
__host__ __device__ inline int choose_k_splits(int tiles, int workers, int max_splits, int min_splits = 1) {
  int s = min_splits > 1 ? min_splits : 1;
  while (s < max_splits) {
    const long long units = static_cast<long long>(tiles) * s;
    const long long rounds = (units + workers - 1) / workers;
    if (units * 100 >= rounds * workers * 95) break;
    s *= 2;
  }
  return s;
}

// The work-unit schedule of one launch.  A pure function of its arguments: the kernels, the host (scratch sizing, the
// launch of the tail-combine kernel) and the combine kernels all derive it from the same values.
constexpr int MAX_TAIL_SPLITS = 16;
struct UnitPlan {
  int num_tiles, num_kb, splits, ks_major;
  int full_tiles;   // tiles that run as whole units (== num_tiles when the tail is not split)
  int tail_splits;  // K pieces per tail tile (1 = no tail splitting)
  int num_units;
};
__host__ __device__ inline UnitPlan make_unit_plan(int num_tiles, int workers, int num_kb, int max_k_splits,
                                                   int min_k_splits, int ks_major, int tail_split) {
  UnitPlan p;
  p.num_tiles = num_tiles;
  p.num_kb = num_kb;
  p.splits = choose_k_splits(num_tiles, workers, max_k_splits > 0 ? max_k_splits : 1, min_k_splits);
  p.ks_major = ks_major;
  p.full_tiles = num_tiles;
  p.tail_splits = 1;
  const int tail = workers > 0 ? num_tiles % workers : 0;
  if (tail_split && p.splits == 1 && tail > 0 && num_tiles > workers && workers / tail >= 2) {
    p.tail_splits = workers / tail < MAX_TAIL_SPLITS ? workers / tail : MAX_TAIL_SPLITS;
    p.full_tiles = num_tiles - tail;
  }
  p.num_units = p.tail_splits > 1 ? p.full_tiles + tail * p.tail_splits : num_tiles * p.splits;
  return p;
}

// Work unit -> (tile, k-split) and the unit's k-block range [kb0, kb1).
struct UnitRange {
  int tile, ks, kb0, kb1, tail_slot;
};
__device__ __forceinline__ UnitRange unit_range(const UnitPlan& p, int unit) {
  UnitRange u;
  int pieces = p.splits;
  u.tail_slot = -1;
  if (p.tail_splits > 1) {
    if (unit < p.full_tiles) {
      u.tile = unit;
      u.ks = 0;
      pieces = 1;
    } else {
      u.tail_slot = unit - p.full_tiles;
      u.tile = p.full_tiles + u.tail_slot / p.tail_splits;
      u.ks = u.tail_slot - (u.tile - p.full_tiles) * p.tail_splits;
      pieces = p.tail_splits;
    }
  } else if (p.ks_major) {
    u.ks = unit / p.num_tiles;
    u.tile = unit - u.ks * p.num_tiles;
  } else {
    u.tile = unit / p.splits;
    u.ks = unit - u.tile * p.splits;
  }
  const int per = (p.num_kb + pieces - 1) / pieces;
  u.kb0 = u.ks * per;
  u.kb1 = u.kb0 + per < p.num_kb ? u.kb0 + per : p.num_kb;
  if (u.kb0 > p.num_kb) u.kb0 = p.num_kb;
  if (u.kb1 < u.kb0) u.kb1 = u.kb0;
  return u;
}

__device__ __forceinline__ TileCoord tile_coord(const GemmShape& s, int tile) {
  TileCoord c;
  if (s.m_fastest) {
    const int per_group = s.group_m * s.tiles_n;
    const int g = tile / per_group;
    const int r = tile - g * per_group;
    const int left = s.tiles_m - g * s.group_m;
    const int gm = left < s.group_m ? left : s.group_m;  // the last group may be short
    c.tile_m = g * s.group_m + r % gm;
    c.tile_n = r / gm;
  } else {
    c.tile_n = tile % s.tiles_n;
    c.tile_m = tile / s.tiles_n;
  }
  c.m0 = c.tile_m * BLOCK_M;
  c.n0 = c.tile_n * BLOCK_N;
  c.ks = 0;
  c.tail_slot = -1;
  c.row_off = 0;
  return c;
}

// Epilogue contract:
//   struct Epi { struct Params {...};
//     __device__ Epi(const Params&, const GemmShape&, uint8_t* warp_smem /*EPI_WARP_SMEM_BYTES, 1024-aligned*/);
//     // called by each of the 128 epilogue threads; `row_in_tile` = TMEM lane owned by the thread
//     __device__ void tile_begin(const TileCoord&, int row_in_tile);
//     __device__ void chunk(const TileCoord&, int row_in_tile, int col_in_tile /*0,32,..*/, const uint32_t (&acc)[32]);
//     __device__ void tile_end(const TileCoord&, int row_in_tile);
//     __device__ void finish();   // once per thread after the last tile (drain outstanding TMA stores)
//   };
// Params is a __grid_constant__ kernel parameter, so it may embed CUtensorMaps for TMA stores.
template <class Epi, bool A_MN, bool B_MN>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
    gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                        const GemmShape shape_in, const __grid_constant__ typename Epi::Params epi_params) {
  GemmShape shape = shape_in;
  if (shape.dyn_rows != nullptr) {
    int r = __ldg(shape.dyn_rows);
    if (shape.dyn_is_k) {
      shape.K = r < shape.K ? (r < 0 ? 0 : r) : shape.K;
    } else {
      shape.M = r < shape.M ? (r < 0 ? 0 : r) : shape.M;
      shape.tiles_m = (shape.M + BLOCK_M - 1) / BLOCK_M;
      if (shape.group_m > shape.tiles_m) shape.group_m = shape.tiles_m > 0 ? shape.tiles_m : 1;
    }
  }
  extern __shared__ uint8_t smem_raw[];
  // 128B-swizzled tiles must be 1024-byte aligned (descriptor base_offset = 0).
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + NUM_STAGES * A_STAGE_BYTES;
  uint8_t* smem_epi = smem + NUM_STAGES * STAGE_BYTES;  // 1024-aligned: NUM_STAGES * STAGE_BYTES is a multiple of 1024
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + NUM_STAGES * STAGE_BYTES + EPI_SMEM_BYTES);
  uint64_t* full_bar = bars;                                    // [NUM_STAGES]  TMA -> MMA
  uint64_t* empty_bar = bars + NUM_STAGES;                      // [NUM_STAGES]  MMA -> TMA
  uint64_t* acc_full_bar = bars + 2 * NUM_STAGES;               // [2] MMA -> epilogue
  uint64_t* acc_empty_bar = bars + 2 * NUM_STAGES + NUM_ACC_STAGES;  // [2] epilogue -> MMA
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 2 * NUM_STAGES + 2 * NUM_ACC_STAGES);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_tiles = shape.tiles_m * shape.tiles_n;
  const int num_kb = (shape.K + BLOCK_K - 1) / BLOCK_K;
  const UnitPlan plan = make_unit_plan(num_tiles, gridDim.x, num_kb, shape.max_k_splits, shape.min_k_splits, shape.ks_major,
                                       shape.tail_split && !shape.m_fastest && (shape.dyn_rows == nullptr || shape.dyn_is_k));
  const int num_units = plan.num_units;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    for (int i = 0; i < NUM_STAGES; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < NUM_ACC_STAGES; ++i) {
      mbar_init(&acc_full_bar[i], 1);
      mbar_init(&acc_empty_bar[i], NUM_EPI_WARPS);
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<1>(tmem_ptr_smem, TMEM_COLS);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 0) {
    // ------------------------------- TMA producer -------------------------------
    if (lane == 0) {
      const uint64_t pol_a = shape.a_l2_policy ? shape.a_l2_policy : L2_EVICT_NORMAL;
      const uint64_t pol_b = shape.b_l2_policy ? shape.b_l2_policy : L2_EVICT_NORMAL;
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int unit = blockIdx.x; unit < num_units; unit += gridDim.x, ++it) {
        const UnitRange ur = unit_range(plan, unit);
        const TileCoord tc = tile_coord(shape, ur.tile);
        const bool rev = shape.k_pingpong && (it & 1);
        for (int ki = ur.kb0; ki < ur.kb1; ++ki) {
          const int kb = rev ? ur.kb0 + ur.kb1 - 1 - ki : ki;
          mbar_wait(&empty_bar[stage], phase ^ 1);
          mbar_arrive_expect_tx(&full_bar[stage], STAGE_BYTES);
          uint8_t* sa = smem_a + stage * A_STAGE_BYTES;
          uint8_t* sb = smem_b + stage * B_STAGE_BYTES;
          const int k0 = kb * BLOCK_K;
          if constexpr (!A_MN) {
            tma_load_2d(&tmap_a, &full_bar[stage], sa, k0, tc.m0, pol_a);  // box {64 k, 128 rows}
          } else {
#pragma unroll
            for (int j = 0; j < BLOCK_M / 64; ++j)  // box {64 m, 64 k-rows} per 64-wide M block
              tma_load_2d(&tmap_a, &full_bar[stage], sa + j * (64 * BLOCK_K * 2), tc.m0 + j * 64, k0, pol_a);
          }
          if constexpr (!B_MN) {
            tma_load_2d(&tmap_b, &full_bar[stage], sb, k0, tc.n0, pol_b);  // box {64 k, 256 rows}
          } else {
#pragma unroll
            for (int j = 0; j < BLOCK_N / 64; ++j)
              tma_load_2d(&tmap_b, &full_bar[stage], sb + j * (64 * BLOCK_K * 2), tc.n0 + j * 64, k0, pol_b);
          }
          if (++stage == NUM_STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------- MMA issuer -------------------------------
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(BLOCK_M, BLOCK_N, A_MN ? 1 : 0, B_MN ? 1 : 0);
      // K-major, 128B swizzle: rows of 128 B, 8-row groups 1024 B apart (SBO); LBO is ignored (16 B, as CuTe encodes it).
      // MN-major, 128B swizzle: each 64-wide MN block is [BLOCK_K rows][128 B]; 8 k-rows per
      // 1024 B group (SBO), next MN block BLOCK_K*128 B further (LBO).
      constexpr uint32_t A_LBO = A_MN ? BLOCK_K * 128 : 16, B_LBO = B_MN ? BLOCK_K * 128 : 16;
      constexpr uint32_t A_KSTEP = A_MN ? UMMA_K * 128 : UMMA_K * 2, B_KSTEP = B_MN ? UMMA_K * 128 : UMMA_K * 2;
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int unit = blockIdx.x; unit < num_units; unit += gridDim.x, ++it) {
        const UnitRange ur = unit_range(plan, unit);
        const int as = it & 1;
        const uint32_t aphase = (it >> 1) & 1;
        mbar_wait(&acc_empty_bar[as], aphase ^ 1);
        tc_fence_after_sync();
        const uint32_t d_tmem = tmem_base + as * BLOCK_N;
        for (int kb = ur.kb0; kb < ur.kb1; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after_sync();
          const uint32_t a_addr = smem_u32(smem_a + stage * A_STAGE_BYTES);
          const uint32_t b_addr = smem_u32(smem_b + stage * B_STAGE_BYTES);
#pragma unroll
          for (int kk = 0; kk < BLOCK_K / UMMA_K; ++kk) {
            const uint64_t da = make_smem_desc_sw128(a_addr + kk * A_KSTEP, A_LBO, 1024);
            const uint64_t db = make_smem_desc_sw128(b_addr + kk * B_KSTEP, B_LBO, 1024);
            umma_bf16<1>(d_tmem, da, db, idesc, (kb != ur.kb0 || kk != 0) ? 1u : 0u);
          }
          umma_commit<1>(&empty_bar[stage]);  // frees the smem slot once these MMAs retire
          if (++stage == NUM_STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit<1>(&acc_full_bar[as]);  // accumulator complete -> epilogue
      }
    }
    __syncwarp();
  } else {
    // ------------------------------- epilogue warps -------------------------------
    const int quarter = warp & 3;  // TMEM lane quarter this warp may access
    const int row_in_tile = quarter * 32 + lane;
    Epi epi(epi_params, shape, smem_epi + (warp - 2) * EPI_WARP_SMEM_BYTES);
    int it = 0;
    for (int unit = blockIdx.x; unit < num_units; unit += gridDim.x, ++it) {
      const UnitRange ur = unit_range(plan, unit);
      TileCoord tc = tile_coord(shape, ur.tile);
      tc.ks = ur.ks;
      tc.tail_slot = ur.tail_slot;
      const int as = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      epi.tile_begin(tc, row_in_tile);
      mbar_wait(&acc_full_bar[as], aphase);
      tc_fence_after_sync();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + as * BLOCK_N;
#pragma unroll 1
      for (int c = 0; c < BLOCK_N; c += 32) {
        uint32_t acc[32];
        __syncwarp();  // epilogue functors may diverge per row; tcgen05.ld is warp-collective
        if (ur.kb1 > ur.kb0) {
          tmem_ld_32x32b_x32(taddr + c, acc);
          tmem_ld_wait();
        } else {  // empty contraction (no live rows): the accumulator was never written
#pragma unroll
          for (int j = 0; j < 32; ++j) acc[j] = 0u;
        }
        epi.chunk(tc, row_in_tile, c, acc);
      }
      tc_fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty_bar[as]);
      epi.tile_end(tc, row_in_tile);
    }
    epi.finish();
  }

  tc_fence_before_sync();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after_sync();
    tmem_dealloc<1>(tmem_base, TMEM_COLS);
  }
}


// =============================================================================================
// cta_group::2 variant: a cluster of two CTAs (an SM pair) works on one 256 x 256 output tile.
//
// Each CTA streams its own 128 rows of A and its own half (128 rows) of B -- 32 KiB per k-block
// instead of 48, so the ring is 6 deep -- and the leader CTA's MMA thread issues
// tcgen05.mma.cta_group::2 (M = 256), which reads A and B from both CTAs' shared memory and
// accumulates 128 rows in each CTA's TMEM.  TMA completions of both CTAs are credited to the
// leader's "full" barrier; tcgen05.commit multicasts to both CTAs' "empty" / "accumulator full"
// barriers; both CTAs' epilogue warps arrive on the leader's "accumulator empty" barrier.
// Same epilogue functors; TileCoord::m0 is this CTA's first row.
// =============================================================================================
namespace pair {

constexpr int TILE_M = 2 * BLOCK_M;             // 256 rows per CTA pair
constexpr int A_BYTES = BLOCK_M * BLOCK_K * 2;  // 16 KiB

// Tile width BN = 256 (used everywhere) or 128 (kept for experiments: it halves wave quantisation for
// GEMMs with few long tiles, but costs 1.5x the operand traffic per flop and measured slower).
template <int BN>
struct Cfg {
  static_assert(BN == 256 || BN == 128, "pair tile width");
  static constexpr int B_ROWS_PER_CTA = BN / 2;
  static constexpr int B_BYTES = B_ROWS_PER_CTA * BLOCK_K * 2;
  static constexpr int CTA_STAGE_BYTES = A_BYTES + B_BYTES;  // 32 / 24 KiB per CTA per stage
  static constexpr int STAGES = (BN == 256) ? 6 : 8;
  static constexpr int TMEM_COLS = NUM_ACC_STAGES * BN;       // 512 / 256
  static constexpr int SMEM_BYTES = STAGES * CTA_STAGE_BYTES + EPI_SMEM_BYTES + 1024 + SMEM_BARRIER_BYTES;
};

template <int BN>
__device__ __forceinline__ TileCoord tile_coord(const GemmShape& s, int tile, int cta_rank) {
  TileCoord c;  // s.tiles_m / s.group_m count 256-row tiles here
  if (s.m_fastest) {
    const int per_group = s.group_m * s.tiles_n;
    const int g = tile / per_group;
    const int r = tile - g * per_group;
    const int left = s.tiles_m - g * s.group_m;
    const int gm = left < s.group_m ? left : s.group_m;
    c.tile_m = g * s.group_m + r % gm;
    c.tile_n = r / gm;
  } else {
    c.tile_n = tile % s.tiles_n;
    c.tile_m = tile / s.tiles_n;
  }
  c.m0 = c.tile_m * TILE_M + cta_rank * BLOCK_M;
  c.n0 = c.tile_n * BN;
  c.ks = 0;
  c.tail_slot = -1;
  c.row_off = cta_rank * BLOCK_M;
  return c;
}

}  // namespace pair

template <class Epi, bool A_MN, bool B_MN, int BN = 256>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
    gemm_tcgen05_2cta_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                             const GemmShape shape_in, const __grid_constant__ typename Epi::Params epi_params) {
  using namespace pair;
  using C = Cfg<BN>;
  constexpr int STAGES = C::STAGES, B_BYTES = C::B_BYTES, CTA_STAGE_BYTES = C::CTA_STAGE_BYTES,
                B_ROWS_PER_CTA = C::B_ROWS_PER_CTA, TMEM_COLS = C::TMEM_COLS;
  GemmShape shape = shape_in;  // tiles_m / group_m arrive in units of 256-row tiles
  if (shape.dyn_rows != nullptr) {
    int r = __ldg(shape.dyn_rows);
    if (shape.dyn_is_k) {
      shape.K = r < shape.K ? (r < 0 ? 0 : r) : shape.K;
    } else {
      shape.M = r < shape.M ? (r < 0 ? 0 : r) : shape.M;
      shape.tiles_m = (shape.M + TILE_M - 1) / TILE_M;
      if (shape.group_m > shape.tiles_m) shape.group_m = shape.tiles_m > 0 ? shape.tiles_m : 1;
    }
  }
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + STAGES * A_BYTES;
  uint8_t* smem_epi = smem + STAGES * CTA_STAGE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * CTA_STAGE_BYTES + EPI_SMEM_BYTES);
  uint64_t* full_bar = bars;                                  // [STAGES]  used in the leader CTA
  uint64_t* empty_bar = bars + STAGES;                        // [STAGES]  per CTA (multicast commit)
  uint64_t* acc_full_bar = bars + 2 * STAGES;                 // [2]       per CTA (multicast commit)
  uint64_t* acc_empty_bar = bars + 2 * STAGES + NUM_ACC_STAGES;  // [2]    used in the leader CTA
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 2 * NUM_ACC_STAGES);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int cta_rank = static_cast<int>(cluster_ctarank());
  const int pair_id = blockIdx.x >> 1;
  const int num_pairs = gridDim.x >> 1;
  const int num_tiles = shape.tiles_m * shape.tiles_n;
  const int num_kb = (shape.K + BLOCK_K - 1) / BLOCK_K;
  const UnitPlan plan = make_unit_plan(num_tiles, num_pairs, num_kb, shape.max_k_splits, shape.min_k_splits, shape.ks_major,
                                       shape.tail_split && !shape.m_fastest && (shape.dyn_rows == nullptr || shape.dyn_is_k));
  const int num_units = plan.num_units;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < NUM_ACC_STAGES; ++i) {
      mbar_init(&acc_full_bar[i], 1);
      mbar_init(&acc_empty_bar[i], 2 * NUM_EPI_WARPS);  // epilogue warps of both CTAs
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<2>(tmem_ptr_smem, TMEM_COLS);
  tc_fence_before_sync();
  cluster_sync_all();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 0) {
    // ------------------------------- TMA producer (both CTAs) -------------------------------
    if (lane == 0) {
      const uint64_t pol_a = shape.a_l2_policy ? shape.a_l2_policy : L2_EVICT_NORMAL;
      const uint64_t pol_b = shape.b_l2_policy ? shape.b_l2_policy : L2_EVICT_NORMAL;
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int unit = pair_id; unit < num_units; unit += num_pairs, ++it) {
        const UnitRange ur = unit_range(plan, unit);
        const TileCoord tc = tile_coord<BN>(shape, ur.tile, cta_rank);
        const int b_row0 = tc.n0 + cta_rank * B_ROWS_PER_CTA;
        const bool rev = shape.k_pingpong && (it & 1);
        for (int ki = ur.kb0; ki < ur.kb1; ++ki) {
          const int kb = rev ? ur.kb0 + ur.kb1 - 1 - ki : ki;
          mbar_wait(&empty_bar[stage], phase ^ 1);
          if (cta_rank == 0) mbar_arrive_expect_tx(&full_bar[stage], 2 * CTA_STAGE_BYTES);  // both CTAs' bytes
          uint8_t* sa = smem_a + stage * A_BYTES;
          uint8_t* sb = smem_b + stage * B_BYTES;
          const int k0 = kb * BLOCK_K;
          if constexpr (!A_MN) {
            tma_load_2d_pair(&tmap_a, &full_bar[stage], sa, k0, tc.m0, pol_a);  // box {64 k, 128 rows}
          } else {
#pragma unroll
            for (int j = 0; j < BLOCK_M / 64; ++j)
              tma_load_2d_pair(&tmap_a, &full_bar[stage], sa + j * (64 * BLOCK_K * 2), tc.m0 + j * 64, k0, pol_a);
          }
          if constexpr (!B_MN) {
            tma_load_2d_pair(&tmap_b, &full_bar[stage], sb, k0, b_row0, pol_b);  // box {64 k, 128 rows}
          } else {
#pragma unroll
            for (int j = 0; j < B_ROWS_PER_CTA / 64; ++j)
              tma_load_2d_pair(&tmap_b, &full_bar[stage], sb + j * (64 * BLOCK_K * 2), b_row0 + j * 64, k0, pol_b);
          }
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------- MMA issuer (leader CTA only) -------------------------------
    if (cta_rank == 0 && lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(TILE_M, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
      constexpr uint32_t A_LBO = A_MN ? BLOCK_K * 128 : 16, B_LBO = B_MN ? BLOCK_K * 128 : 16;
      constexpr uint32_t A_KSTEP = A_MN ? UMMA_K * 128 : UMMA_K * 2, B_KSTEP = B_MN ? UMMA_K * 128 : UMMA_K * 2;
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int unit = pair_id; unit < num_units; unit += num_pairs, ++it) {
        const UnitRange ur = unit_range(plan, unit);
        const int as = it & 1;
        const uint32_t aphase = (it >> 1) & 1;
        mbar_wait(&acc_empty_bar[as], aphase ^ 1);
        tc_fence_after_sync();
        const uint32_t d_tmem = tmem_base + as * BN;
        for (int kb = ur.kb0; kb < ur.kb1; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after_sync();
          const uint32_t a_addr = smem_u32(smem_a + stage * A_BYTES);
          const uint32_t b_addr = smem_u32(smem_b + stage * B_BYTES);
#pragma unroll
          for (int kk = 0; kk < BLOCK_K / UMMA_K; ++kk) {
            const uint64_t da = make_smem_desc_sw128(a_addr + kk * A_KSTEP, A_LBO, 1024);
            const uint64_t db = make_smem_desc_sw128(b_addr + kk * B_KSTEP, B_LBO, 1024);
            umma_bf16<2>(d_tmem, da, db, idesc, (kb != ur.kb0 || kk != 0) ? 1u : 0u);
          }
          umma_commit<2>(&empty_bar[stage], 0x3);  // frees the stage in both CTAs
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit<2>(&acc_full_bar[as], 0x3);  // accumulator halves complete in both CTAs
      }
    }
    __syncwarp();
  } else {
    // ------------------------------- epilogue warps (both CTAs) -------------------------------
    const int quarter = warp & 3;
    const int row_in_tile = quarter * 32 + lane;  // row within this CTA's 128-row half
    Epi epi(epi_params, shape, smem_epi + (warp - 2) * EPI_WARP_SMEM_BYTES);
    int it = 0;
    for (int unit = pair_id; unit < num_units; unit += num_pairs, ++it) {
      const UnitRange ur = unit_range(plan, unit);
      TileCoord tc = tile_coord<BN>(shape, ur.tile, cta_rank);
      tc.ks = ur.ks;
      tc.tail_slot = ur.tail_slot;
      const int as = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      epi.tile_begin(tc, row_in_tile);
      mbar_wait(&acc_full_bar[as], aphase);
      tc_fence_after_sync();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + as * BN;
#pragma unroll 1
      for (int c = 0; c < BN; c += 32) {
        uint32_t acc[32];
        __syncwarp();
        if (ur.kb1 > ur.kb0) {
          tmem_ld_32x32b_x32(taddr + c, acc);
          tmem_ld_wait();
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) acc[j] = 0u;
        }
        epi.chunk(tc, row_in_tile, c, acc);
      }
      tc_fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(&acc_empty_bar[as], 0);  // the leader's barrier
      epi.tile_end(tc, row_in_tile);
    }
    epi.finish();
  }

  tc_fence_before_sync();
  cluster_sync_all();  // neither CTA may exit (or free TMEM) while its peer can still touch its smem / barriers
  if (warp == 1) {
    tc_fence_after_sync();
    tmem_dealloc<2>(tmem_base, TMEM_COLS);
  }
}

}  // namespace b200
