// tcgen05 / TMEM GEMM fed by TMA: D[M,N] = alpha * A[M,K] · B[N,K]^T (+ bias, dropout, residual).
//
// Replaces the cuBLAS/cuDNN calls behind every nn.Linear of the decoder (K10, K11, K13) and every 1x1
// convolution of the image tokenizer (K3, K6, K8) — forward, data-gradient and weight-gradient alike, which is
// why both operands may be K-major *or* MN-major (a transposed view costs nothing: the transposition happens
// in the TMA box shape and the UMMA descriptor, never in memory).
//
// Structure (one 128 x BLOCK_N output tile per CTA, 128 threads, non-persistent — several CTAs are resident
// per SM and overlap each other's load / MMA / epilogue phases, which suits the short-K, memory-bound shapes
// of this model):
//   warp 0 / lane 0 : TMA producer   — cp.async.bulk.tensor 2-D boxes (128B swizzle) into a ring of stages,
//                                       completion counted on "full" mbarriers
//   warp 1 / lane 0 : MMA issuer     — 4 x tcgen05.mma (K = 16) per 64-wide k-block, accumulator in TMEM,
//                                       tcgen05.commit releases the stage ("empty") and finally signals the
//                                       epilogue ("tmem_full")
//   warps 0-3       : epilogue       — tcgen05.ld 32x32b (thread ↔ accumulator row), fused bias / dropout /
//                                       residual / scale, bf16 or fp32 stores, or fp32 atomics for split-K
// Partial tiles need no special code on the load side: TMA zero-fills out-of-bounds box elements.
#include <stdlib.h>

#include "sm100.cuh"
#include "bn_tail.cuh"

using namespace sm100;

struct GemmParams {
  int M, N, K;
  int a_mn, b_mn;        // 1 = operand stored MN-major (i.e. as [K, M] / [K, N] row-major)
  int block_n;           // multiple of 16, <= 256
  int stages;
  int kb_per_split;      // 64-wide k-blocks handled by one CTA (grid.z splits)
  void* out;
  long ldo;
  int out_fp32;
  int atomic_out;        // fp32 atomicAdd into `out` (split-K / accumulate)
  const bf16* bias;      // [N] or null
  const bf16* residual;  // [M, ldr] or null
  long ldr;
  float alpha;
  float dropout_p;
  uint32_t seed;
  const uint32_t* seed_epoch;   // device word added to `seed` (or null), see rt1_seed_epoch_ptr()
  // optional per-column statistics of the (post-bias) fp32 output: sum and sum of squares, atomically added
  float* col_sum;
  float* col_sumsq;
  int b_resident;        // persistent kernel: all B tiles stay in shared memory (small weights, huge M)
  int has_tail;          // persistent kernel: the CTA that lands the last statistics finalises the BatchNorm (bn_tail.cuh)
  Rt1BnTail tail;
  int dbg;               // bench-only switches (RT1B200_GEMM_DBG): 1 no copy-out, 2 no statistics, 4 no TMEM drain/staging, 8 no A loads, 16 no B loads
};

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;
constexpr int A_TILE_BYTES = BLOCK_M * BLOCK_K * 2;  // 16 KiB

__host__ __device__ inline int b_tile_bytes(int block_n, int b_mn) {
  return b_mn ? ((block_n + 63) / 64) * (BLOCK_K * 128) : block_n * 128;
}

__global__ void __launch_bounds__(128, 8) gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tma_a,
                                                           const __grid_constant__ CUtensorMap tma_b, const GemmParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // dynamic smem base is only guaranteed 16B-aligned by the ABI: align up to 1024 (swizzle-128B atom)
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // pointer arithmetic keeps the shared address space (LDS/STS, not generic LD/ST)
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // 1-D grid over (m-tile, n-tile) with the n-tile index fastest: the CTAs that share an A tile run back to back,
  // so A is fetched from HBM once and re-read from L2
  const int n_tiles = (p.N + p.block_n - 1) / p.block_n;
  const int m0 = (blockIdx.x / n_tiles) * BLOCK_M, n0 = (blockIdx.x % n_tiles) * p.block_n;
  const int total_kb = (p.K + BLOCK_K - 1) / BLOCK_K;
  const int kb_begin = blockIdx.z * p.kb_per_split;
  const int kb_end = min(total_kb, kb_begin + p.kb_per_split);
  const int num_kb = kb_end - kb_begin;
  const int b_bytes = b_tile_bytes(p.block_n, p.b_mn);
  const int stage_bytes = A_TILE_BYTES + b_bytes;

  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)p.stages * stage_bytes);
  const uint32_t full0 = smem_u32(bars), empty0 = smem_u32(bars + p.stages), tmem_full = smem_u32(bars + 2 * p.stages);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * p.stages + 1);
  const uint32_t smem_base = smem_u32(smem);

  uint32_t tmem_cols = 32;
  while ((int)tmem_cols < p.block_n) tmem_cols <<= 1;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(full0 + 8 * s, 1);
      mbar_init(empty0 + 8 * s, 1);
    }
    mbar_init(tmem_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(smem_u32(tmem_slot), tmem_cols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_enter();

  if (num_kb > 0) {
    if (warp == 0 && lane == 0) {
      // ------------------------------------------------------------------ TMA producer
      for (int i = 0; i < num_kb; ++i) {
        const int s = i % p.stages;
        const uint32_t ph = (uint32_t)(i / p.stages) & 1u;
        mbar_wait(empty0 + 8 * s, ph ^ 1u);
        const uint32_t a_dst = smem_base + s * stage_bytes, b_dst = a_dst + A_TILE_BYTES;
        const uint32_t bar = full0 + 8 * s;
        mbar_expect_tx(bar, (uint32_t)stage_bytes);
        const int kg = (kb_begin + i) * BLOCK_K;
        if (!p.a_mn) {
          tma_load_2d(a_dst, &tma_a, bar, kg, m0);
        } else {
          tma_load_2d(a_dst, &tma_a, bar, m0, kg);
          tma_load_2d(a_dst + BLOCK_K * 128, &tma_a, bar, m0 + 64, kg);
        }
        if (!p.b_mn) {
          tma_load_2d(b_dst, &tma_b, bar, kg, n0);
        } else {
          for (int j = 0; j * 64 < p.block_n; ++j) tma_load_2d(b_dst + j * (BLOCK_K * 128), &tma_b, bar, n0 + 64 * j, kg);
        }
      }
    } else if (warp == 1 && lane == 0) {
      // ------------------------------------------------------------------ MMA issuer
      const uint32_t idesc = umma_idesc_bf16(BLOCK_M, p.block_n, p.a_mn, p.b_mn);
      for (int i = 0; i < num_kb; ++i) {
        const int s = i % p.stages;
        const uint32_t ph = (uint32_t)(i / p.stages) & 1u;
        mbar_wait(full0 + 8 * s, ph);
        tc_fence_after();
        const uint32_t a_src = smem_base + s * stage_bytes, b_src = a_src + A_TILE_BYTES;
#pragma unroll
        for (int k = 0; k < BLOCK_K / 16; ++k) {
          const uint64_t da = p.a_mn ? umma_smem_desc(a_src + k * 2048, BLOCK_K * 128, 1024)
                                     : umma_smem_desc(a_src + k * 32, 16, 1024);
          const uint64_t db = p.b_mn ? umma_smem_desc(b_src + k * 2048, BLOCK_K * 128, 1024)
                                     : umma_smem_desc(b_src + k * 32, 16, 1024);
          umma_bf16(tmem_base, da, db, idesc, (i | k) != 0);
        }
        umma_commit(empty0 + 8 * s);               // stage reusable once these MMAs have read it
        if (i == num_kb - 1) umma_commit(tmem_full);  // accumulator complete
      }
    }
    __syncwarp();
    // -------------------------------------------------------------------- epilogue (all 4 warps)
    mbar_wait(tmem_full, 0);
    tc_fence_after();
  }

  const int row = m0 + warp * 32 + lane;
  const bool row_ok = row < p.M;
  const uint32_t thresh = rt1_drop_threshold(p.dropout_p);
  const float keep_scale = p.dropout_p > 0.f ? 1.f / (1.f - p.dropout_p) : 1.f;
  float* stat_sm = reinterpret_cast<float*>(smem);  // operand stages are dead once tmem_full fired

  // 16 accumulator columns at a time: keeps the kernel at <= 64 registers/thread so 8 CTAs fit on an SM
  // (with 32-column chunks ptxas used 128 registers → 4 CTAs/SM and the epilogue latency was exposed).
  for (int c0 = 0; c0 < p.block_n; c0 += 16) {
    uint32_t acc[16];
    if (num_kb > 0) {
      tmem_ld_32x16(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, acc);
      tmem_ld_wait();
    } else {
#pragma unroll
      for (int j = 0; j < 16; ++j) acc[j] = 0u;
    }
    float v[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const int col = n0 + c0 + j;
      float x = __uint_as_float(acc[j]) * p.alpha;
      if (p.bias != nullptr && col < p.N) x += __bfloat162float(p.bias[col]);
      v[j] = x;
    }
    if (p.col_sum != nullptr) {
      // per-column sum / sum-of-squares over the 128 rows of this tile: transpose through shared memory
      // (17-float pitch: conflict-free), 8 row-groups x 16 columns reduce in parallel, then 16 threads fold the
      // partials and issue one atomic pair per column per CTA
      __syncthreads();
#pragma unroll
      for (int j = 0; j < 16; ++j) stat_sm[threadIdx.x * 17 + j] = row_ok ? v[j] : 0.f;
      __syncthreads();
      {
        const int col_l = threadIdx.x & 15, grp = threadIdx.x >> 4;  // 8 groups of 16 rows
        float s = 0.f, q = 0.f;
#pragma unroll
        for (int r = 0; r < 16; ++r) {
          const float t = stat_sm[(grp * 16 + r) * 17 + col_l];
          s += t;
          q = fmaf(t, t, q);
        }
        float* part = stat_sm + 128 * 17;
        part[grp * 32 + col_l] = s;
        part[grp * 32 + 16 + col_l] = q;
        __syncthreads();
        if (threadIdx.x < 16) {
          s = q = 0.f;
#pragma unroll
          for (int g8 = 0; g8 < 8; ++g8) {
            s += part[g8 * 32 + threadIdx.x];
            q += part[g8 * 32 + 16 + threadIdx.x];
          }
          const int col = n0 + c0 + threadIdx.x;
          if (col < p.N) {
            atomicAdd(p.col_sum + col, s);
            atomicAdd(p.col_sumsq + col, q);
          }
        }
      }
    }
    if (!row_ok) continue;
    if (p.dropout_p > 0.f) {
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const uint32_t idx = (uint32_t)((long)row * p.N + (n0 + c0 + j));
        v[j] = rt1_hash_u32(idx, p.seed + (p.seed_epoch != nullptr ? __ldg(p.seed_epoch) : 0u)) >= thresh ? v[j] * keep_scale : 0.f;
      }
    }
    if (p.residual != nullptr) {
      const bf16* rr = p.residual + (long)row * p.ldr + n0 + c0;
#pragma unroll
      for (int j8 = 0; j8 < 16; j8 += 8) {
        const int col = n0 + c0 + j8;
        if (col + 8 <= p.N && ((reinterpret_cast<uintptr_t>(rr + j8) & 15) == 0)) {
          float r8[8];
          unpack8(*reinterpret_cast<const bf16x8*>(rr + j8), r8);
#pragma unroll
          for (int k = 0; k < 8; ++k) v[j8 + k] += r8[k];
        } else {
#pragma unroll
          for (int k = 0; k < 8; ++k)
            if (col + k < p.N) v[j8 + k] += __bfloat162float(rr[j8 + k]);
        }
      }
    }
    if (p.out_fp32) {
      float* o = reinterpret_cast<float*>(p.out) + (long)row * p.ldo + n0 + c0;
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        if (n0 + c0 + j < p.N) {
          if (p.atomic_out) atomicAdd(o + j, v[j]);
          else o[j] = v[j];
        }
      }
    } else {
      bf16* o = reinterpret_cast<bf16*>(p.out) + (long)row * p.ldo + n0 + c0;
#pragma unroll
      for (int j8 = 0; j8 < 16; j8 += 8) {
        const int col = n0 + c0 + j8;
        if (col + 8 <= p.N && ((reinterpret_cast<uintptr_t>(o + j8) & 15) == 0)) {
          *reinterpret_cast<bf16x8*>(o + j8) = pack8(v + j8);
        } else {
#pragma unroll
          for (int k = 0; k < 8; ++k)
            if (col + k < p.N) o[j8 + k] = __float2bfloat16_rn(v[j8 + k]);
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, tmem_cols);
}

// =================================================================================================================
// Persistent variant for bf16 outputs (every forward and data-gradient GEMM of the model — the ones that move the
// activation bytes).  The tile-per-CTA kernel above wrote each accumulator row straight from registers: a warp's
// store instruction touched 32 different rows, i.e. 32 partially filled sectors, and ncu showed the 1x1-conv GEMMs at
// 5-8 % of DRAM bandwidth with the LSU/L2 write path saturated.  Here:
//   * one CTA per SM walks the (m-tile, n-tile) list; operand stages form ONE ring across tiles, so the TMA loads of
//     the next tiles are in flight while the current tile is multiplied and the previous ones are written out;
//   * two TMEM accumulators (2 x BLOCK_N columns): the MMA of tile i+1 overlaps the epilogue of tile i;
//   * two epilogue groups of 8 warps (4 TMEM lane quadrants x 2 column halves) alternate tiles: tcgen05.ld → bias /
//     alpha / dropout / residual → bf16 → 128B-swizzled staging tile in shared memory → ONE TMA store per 64-column
//     slab (full 128-byte lines to L2, partial tiles clipped by the tensor map).  16 epilogue warps = 4 per scheduler:
//     with 4 (one per scheduler) every tcgen05.ld / LDS latency was exposed and a 128x128 tile took ~5 us;
//   * BatchNorm column statistics are taken from the staged bf16 tile (the values the consumer will actually read),
//     accumulated per CTA in shared memory and flushed with one atomic per column per CTA at the end.
// Warp roles: 0 = TMA producer, 1 = MMA issuer (+ TMEM alloc), 2-9 = epilogue group 0, 10-17 = epilogue group 1.
constexpr int V2_THREADS = 576;
constexpr int V2_EPI_THREADS = 256;
constexpr int V2_STAGE_TILE_BYTES = BLOCK_M * 128;   // one 64-column bf16 slab of the staging tile (128B rows)

constexpr int V2_MAX_ACC = 4;                         // TMEM accumulator ring (512 columns / 128)
constexpr int V2_BRES_MAX = 64 * 1024;                // B operands up to this size stay resident in shared memory

struct GemmV2Smem {   // byte offsets from the 1024-aligned base
  int stage0, bres, staging, stats, bias, bars, total;
};
__host__ __device__ inline GemmV2Smem v2_smem_layout(int stages, int stage_bytes, int bres_bytes, int n_stats) {
  GemmV2Smem L;
  L.stage0 = 0;
  L.bres = stages * stage_bytes;                             // resident B tiles [n_tile][k_block] (0 bytes in ring mode)
  L.staging = L.bres + bres_bytes;                           // 2 groups x 2 slabs x 16 KiB
  L.stats = L.staging + 2 * 2 * V2_STAGE_TILE_BYTES;         // 2 x n_stats floats
  L.bias = L.stats + 2 * n_stats * 4;                        // 2 groups x 128 floats
  L.bars = (L.bias + 2 * 128 * 4 + 7) & ~7;                  // full[stages], empty[stages], tmem_full[4], tmem_empty[4], bres, slot
  L.total = L.bars + (2 * stages + 2 * V2_MAX_ACC + 1) * 8 + 16;
  return L;
}

// Tile order: n fastest (the n-tiles of one 128-row A slab run back to back on the same SM), serpentine — odd slabs
// walk the n-tiles backwards — so that consecutive tiles, which alternate between the two epilogue groups, alternate
// between wide and narrow n-tiles as well.
__device__ __forceinline__ void v2_tile_coords(int t, int n_tiles, int block_n, int& m0, int& n0) {
  const int mt = t / n_tiles, j = t - mt * n_tiles;
  m0 = mt * BLOCK_M;
  n0 = ((mt & 1) ? n_tiles - 1 - j : j) * block_n;
}

// PLAIN = no alpha / bias / dropout / residual (every 1x1 convolution of the image tokenizer): the epilogue is
// instruction-issue bound, so the four per-8-column feature tests are compiled out for it.
template <bool PLAIN>
__global__ void __launch_bounds__(V2_THREADS, 1) gemm_tcgen05_persistent_kernel(const __grid_constant__ CUtensorMap tma_a,
                                                                              const __grid_constant__ CUtensorMap tma_b,
                                                                              const __grid_constant__ CUtensorMap tma_o,
                                                                              const GemmParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // pointer arithmetic keeps the shared address space (LDS/STS, not generic LD/ST)
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_tiles = (p.N + p.block_n - 1) / p.block_n;
  const int m_tiles = (p.M + BLOCK_M - 1) / BLOCK_M;
  const int num_tiles = n_tiles * m_tiles;
  const int num_kb = (p.K + BLOCK_K - 1) / BLOCK_K;
  // each CTA walks a CONTIGUOUS run of tiles (a strided walk with an even grid would hand one CTA only the wide and
  // its neighbour only the narrow n-tiles of N = 128 + 16)
  const int per_cta = (num_tiles + gridDim.x - 1) / gridDim.x;
  const int t_begin = min(num_tiles, (int)blockIdx.x * per_cta), t_end = min(num_tiles, t_begin + per_cta);
  const int b_bytes = b_tile_bytes(p.block_n, p.b_mn);
  // Weights of the early, huge-M 1x1 convolutions are a few KB: in resident mode every (n-tile, k-block) B tile is
  // loaded ONCE per CTA and the stage ring carries A only.  (Re-loading them per tile cost more than the A stream:
  // boxes that are mostly out of bounds — K = 24 of 64, N = 24 of 64 — are slow in the TMA unit.)
  const int bres_bytes = p.b_resident ? n_tiles * num_kb * b_bytes : 0;
  const int stage_bytes = A_TILE_BYTES + (p.b_resident ? 0 : b_bytes);
  const int n_stats = p.col_sum != nullptr ? ((p.N + 3) & ~3) : 0;
  const GemmV2Smem L = v2_smem_layout(p.stages, stage_bytes, bres_bytes, n_stats);

  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.bars);
  const uint32_t full0 = smem_u32(bars), empty0 = smem_u32(bars + p.stages);
  const uint32_t tfull0 = smem_u32(bars + 2 * p.stages), tempty0 = smem_u32(bars + 2 * p.stages + V2_MAX_ACC);
  const uint32_t bres_bar = smem_u32(bars + 2 * p.stages + 2 * V2_MAX_ACC);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * p.stages + 2 * V2_MAX_ACC + 1);
  const uint32_t smem_base = smem_u32(smem);
  float* s_stats = reinterpret_cast<float*>(smem + L.stats);

  // Accumulator ring: the MMA warp runs up to n_acc tiles ahead of the epilogue, which hides the
  // tempty → MMA issue → commit → tfull handshake chain (~1.4 us end to end, measured with everything else off).
  const int acc_stride = (p.block_n + 31) & ~31;      // the epilogue reads 32 columns at a time
  const uint32_t n_acc = 4 * acc_stride <= 512 ? 4u : 2u;
  uint32_t tmem_cols = 32;
  while (tmem_cols < n_acc * acc_stride) tmem_cols <<= 1;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(full0 + 8 * s, 1);
      mbar_init(empty0 + 8 * s, 1);
    }
    for (int g = 0; g < V2_MAX_ACC; ++g) {
      mbar_init(tfull0 + 8 * g, 1);
      mbar_init(tempty0 + 8 * g, 8);       // one arrival per epilogue warp of the group
    }
    mbar_init(bres_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(smem_u32(tmem_slot), tmem_cols);
    tmem_relinquish();
  }
  for (int i = threadIdx.x; i < 2 * n_stats; i += V2_THREADS) s_stats[i] = 0.f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_enter();

  if (warp == 0) {
    if (lane == 0) {
      // ------------------------------------------------------------------ TMA producer (one ring across tiles)
      uint32_t it = 0;
      if (p.b_resident && t_begin < t_end) {
        mbar_expect_tx(bres_bar, (uint32_t)bres_bytes);
        for (int nt = 0; nt < n_tiles; ++nt)
          for (int kb = 0; kb < num_kb; ++kb) {
            const uint32_t b_dst = smem_base + L.bres + (nt * num_kb + kb) * b_bytes;
            if (!p.b_mn) {
              tma_load_2d(b_dst, &tma_b, bres_bar, kb * BLOCK_K, nt * p.block_n);
            } else {
              for (int j = 0; j * 64 < p.block_n; ++j)
                tma_load_2d(b_dst + j * (BLOCK_K * 128), &tma_b, bres_bar, nt * p.block_n + 64 * j, kb * BLOCK_K);
            }
          }
      }
      uint32_t s = 0, ph = 0;
      int mt = t_begin / n_tiles, j = t_begin - mt * n_tiles;
      for (int t = t_begin; t < t_end; ++t) {
        const int m0 = mt * BLOCK_M, n0 = ((mt & 1) ? n_tiles - 1 - j : j) * p.block_n;
        if (++j == n_tiles) { j = 0; ++mt; }
        for (int kb = 0; kb < num_kb; ++kb, ++it) {
          mbar_wait(empty0 + 8 * s, ph ^ 1u);
          const uint32_t a_dst = smem_base + s * stage_bytes, b_dst = a_dst + A_TILE_BYTES;
          const uint32_t bar = full0 + 8 * s;
          const bool skip_a = (p.dbg & 8) && it >= (uint32_t)p.stages;
          const bool skip_b = p.b_resident || ((p.dbg & 16) && it >= (uint32_t)p.stages);
          mbar_expect_tx(bar, (uint32_t)((skip_a ? 0 : A_TILE_BYTES) + (skip_b ? 0 : b_bytes)));
          const int kg = kb * BLOCK_K;
          if (skip_a) {
          } else if (!p.a_mn) {
            tma_load_2d(a_dst, &tma_a, bar, kg, m0);
          } else {
            tma_load_2d(a_dst, &tma_a, bar, m0, kg);
            tma_load_2d(a_dst + BLOCK_K * 128, &tma_a, bar, m0 + 64, kg);
          }
          if (skip_b) {
          } else if (!p.b_mn) {
            tma_load_2d(b_dst, &tma_b, bar, kg, n0);
          } else {
            for (int jj = 0; jj * 64 < p.block_n; ++jj) tma_load_2d(b_dst + jj * (BLOCK_K * 128), &tma_b, bar, n0 + 64 * jj, kg);
          }
          if (++s == (uint32_t)p.stages) { s = 0; ph ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // ------------------------------------------------------------------ MMA issuer
      // This loop runs on ONE thread and its instruction latency is the pipeline's clock (~0.8 us per tile with
      // divisions and descriptor encoding inside; measured with all memory traffic switched off), so everything is
      // incremental: ring indices and phases are counters, descriptors are a constant plus an address term.
      const uint32_t idesc = umma_idesc_bf16(BLOCK_M, p.block_n, p.a_mn, p.b_mn);
      const uint64_t a_desc0 = p.a_mn ? umma_smem_desc(0, BLOCK_K * 128, 1024) : umma_smem_desc(0, 16, 1024);
      const uint64_t b_desc0 = p.b_mn ? umma_smem_desc(0, BLOCK_K * 128, 1024) : umma_smem_desc(0, 16, 1024);
      const uint32_t a_kstep = p.a_mn ? 2048u >> 4 : 32u >> 4, b_kstep = p.b_mn ? 2048u >> 4 : 32u >> 4;
      uint32_t s = 0, ph = 0, g = 0, gph = 1;                // stage ring index/phase, accumulator ring index/(phase ^ 1)
      int mt = t_begin / n_tiles, j = t_begin - mt * n_tiles; // serpentine walk, see v2_tile_coords
      if (p.b_resident && t_begin < t_end) mbar_wait(bres_bar, 0);
      for (int t = t_begin; t < t_end; ++t) {
        const int nt = (mt & 1) ? n_tiles - 1 - j : j;
        mbar_wait(tempty0 + 8 * g, gph);                     // epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + g * (uint32_t)acc_stride;
        uint32_t b_res = smem_base + L.bres + nt * num_kb * b_bytes;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(full0 + 8 * s, ph);
          tc_fence_after();
          const uint32_t a_src = smem_base + s * stage_bytes;
          const uint32_t b_src = p.b_resident ? b_res : a_src + A_TILE_BYTES;
          const uint64_t da = a_desc0 + (uint64_t)((a_src & 0x3FFFF) >> 4), db = b_desc0 + (uint64_t)((b_src & 0x3FFFF) >> 4);
#pragma unroll
          for (int k = 0; k < BLOCK_K / 16; ++k) umma_bf16(d_tmem, da + k * a_kstep, db + k * b_kstep, idesc, (kb | k) != 0);
          umma_commit(empty0 + 8 * s);
          if (kb == num_kb - 1) umma_commit(tfull0 + 8 * g);
          b_res += b_bytes;
          if (++s == (uint32_t)p.stages) { s = 0; ph ^= 1u; }
        }
        if (++g == n_acc) { g = 0; gph ^= 1u; }
        if (++j == n_tiles) { j = 0; ++mt; }
      }
    }
  } else {
    // -------------------------------------------------------------------- epilogue groups (warps 2-9, 10-17)
    const int grp = (warp - 2) >> 3;                       // 0 / 1: which staging tile this group owns
    const int quad = warp & 3;                             // TMEM lane quadrant this warp may read (warp id % 4)
    const int half = ((warp - 2) & 7) >> 2;                // column half (= staging slab) this warp converts
    const int tid = (((warp - 2) & 7) << 5) | lane;        // 0..255 within the group
    const int row_l = quad * 32 + lane;                    // accumulator row handled by this thread
    uint8_t* stage_g = smem + L.staging + grp * 2 * V2_STAGE_TILE_BYTES;
    float* s_bias = reinterpret_cast<float*>(smem + L.bias) + grp * 128;
    const uint32_t thresh = rt1_drop_threshold(p.dropout_p);
    const float keep_scale = p.dropout_p > 0.f ? 1.f / (1.f - p.dropout_p) : 1.f;
    // Tile cursor (m-tile, position in the serpentine n walk) and accumulator ring position advance by increments: the
    // divisions they replace (t / n_tiles, li % n_acc, li / n_acc) were ~15 % of the epilogue's executed instructions.
    uint32_t li = 0, acc_i = 0, use = 0;
    int mt = t_begin / n_tiles, jn = t_begin - mt * n_tiles;
    for (int t = t_begin; t < t_end; ++t, ++li) {
      const uint32_t acc_cur = acc_i, use_cur = use;
      const int m0 = mt * BLOCK_M, n0 = ((mt & 1) ? n_tiles - 1 - jn : jn) * p.block_n;
      if (++acc_i == n_acc) { acc_i = 0; ++use; }
      if (++jn == n_tiles) { jn = 0; ++mt; }
      if ((int)(li & 1u) != grp) continue;                 // group g owns accumulators g, g + 2: it sees every phase of their barriers
      const int nv = min(p.block_n, p.N - n0);              // valid columns of this tile
      const int row = m0 + row_l;
      const bool row_ok = row < p.M;
      if (tid == 0) tma_store_wait_read();                   // staging tile free again (previous store has read it)
      if (p.bias != nullptr && tid < p.block_n) s_bias[tid] = (n0 + tid < p.N) ? __bfloat162float(p.bias[n0 + tid]) : 0.f;
      named_bar_sync(1 + grp, V2_EPI_THREADS);
      mbar_wait(tfull0 + 8 * acc_cur, use_cur & 1u);
      tc_fence_after();
      const uint32_t t_addr = tmem_base + ((uint32_t)(quad * 32) << 16) + acc_cur * (uint32_t)acc_stride;
      const bool full_tile = m0 + BLOCK_M <= p.M;
      // the accumulator goes back to the MMA warp as soon as its last chunk sits in registers
      const int cbeg = half * 64, cend = (p.dbg & 4) ? 0 : min(nv, half * 64 + 64);
#pragma unroll
      for (int h2 = 0; h2 < 2; ++h2) {
        const int c0 = cbeg + 32 * h2;
        uint32_t acc[32];
        if (c0 < cend) {
          tmem_ld_32x32(t_addr + (uint32_t)c0, acc);
          tmem_ld_wait();
        }
        if (h2 == 1) {                                     // last read of this accumulator: hand it back to the MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(tempty0 + 8 * acc_cur);
        }
        if (c0 >= cend) continue;
        const bool edge = !full_tile || c0 + 32 > nv;
        // rows past M and columns past N are staged as zeros (the copy-out skips them, the statistics ignore them).
        // XOR-swizzled staging: 16-byte chunk c of row r lives at chunk (c ^ (r & 7)) — conflict-free for the
        // row-per-thread writes here and for the 8-lanes-per-row reads of the copy-out below
        uint8_t* slab = stage_g + (c0 >> 6) * V2_STAGE_TILE_BYTES + row_l * 128;
        const int cbase = (c0 & 63) >> 3;
#pragma unroll
        for (int q = 0; q < 4; ++q) {                      // 8 columns at a time keeps the live register set small
          const int c8 = c0 + 8 * q;
          float v[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) v[k] = __uint_as_float(acc[8 * q + k]);
          if (!PLAIN && p.alpha != 1.f) {
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] *= p.alpha;
          }
          if (!PLAIN && p.bias != nullptr) {
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] += s_bias[c8 + k];
          }
          if (!PLAIN && p.dropout_p > 0.f) {
#pragma unroll
            for (int k = 0; k < 8; ++k) {
              const uint32_t idx = (uint32_t)((long)row * p.N + (n0 + c8 + k));
              v[k] = rt1_hash_u32(idx, p.seed + (p.seed_epoch != nullptr ? __ldg(p.seed_epoch) : 0u)) >= thresh ? v[k] * keep_scale : 0.f;
            }
          }
          if (!PLAIN && p.residual != nullptr && row_ok) {
            const bf16* rr = p.residual + (long)row * p.ldr + n0 + c8;
            if (c8 + 8 <= nv && ((reinterpret_cast<uintptr_t>(rr) & 15) == 0)) {
              float r8[8];
              unpack8(*reinterpret_cast<const bf16x8*>(rr), r8);
#pragma unroll
              for (int k = 0; k < 8; ++k) v[k] += r8[k];
            } else {
#pragma unroll
              for (int k = 0; k < 8; ++k)
                if (c8 + k < nv) v[k] += __bfloat162float(rr[k]);
            }
          }
          if (edge) {
#pragma unroll
            for (int k = 0; k < 8; ++k)
              if (!row_ok || c8 + k >= nv) v[k] = 0.f;
          }
          *reinterpret_cast<bf16x8*>(slab + (((cbase + q) ^ (row_l & 7)) << 4)) = pack8(v);
        }
      }
      fence_proxy_async();                                   // staged tile → visible to the TMA (async proxy)
      named_bar_sync(1 + grp, V2_EPI_THREADS);
      // ONE TMA store per 64-column slab: full 128-byte lines to L2, rows past M / columns past N clipped by the
      // tensor map.  (Writing the tile out with ordinary 16-byte stores from the epilogue warps was tried and lost:
      // the epilogue is instruction-issue bound and the copy-out added ~150 instructions per thread and tile.)
      if (tid == 0 && !(p.dbg & 1)) {
        for (int jj = 0; jj * 64 < nv; ++jj)
          tma_store_2d(&tma_o, smem_u32(stage_g) + jj * V2_STAGE_TILE_BYTES, n0 + 64 * jj, m0);
        tma_store_commit();
      }
      if (n_stats > 0 && !(p.dbg & 2)) {
        // column statistics from the staged bf16 values: thread ↔ (column pair, quarter of the rows)
        const int cp = tid & 63, rh = tid >> 6;
        if (2 * cp < nv) {
          const uint8_t* slab = stage_g + (cp >> 5) * V2_STAGE_TILE_BYTES;
          const int chunk = (cp & 31) >> 2, inner = (cp & 3) * 4;
          float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
#pragma unroll 8
          for (int r = rh * 32; r < rh * 32 + 32; ++r) {
            const bf162 x2 = *reinterpret_cast<const bf162*>(slab + r * 128 + ((chunk ^ (r & 7)) << 4) + inner);
            const float2 f = __bfloat1622float2(x2);
            s0 += f.x; q0 = fmaf(f.x, f.x, q0);
            s1 += f.y; q1 = fmaf(f.y, f.y, q1);
          }
          const int col = n0 + 2 * cp;
          atomicAdd(s_stats + col, s0);
          atomicAdd(s_stats + n_stats + col, q0);
          if (col + 1 < p.N) {
            atomicAdd(s_stats + col + 1, s1);
            atomicAdd(s_stats + n_stats + col + 1, q1);
          }
        }
      }
    }
    if (tid == 0) tma_store_wait_read();                     // shared memory must outlive the last bulk store's reads
  }

  tc_fence_before();
  __syncthreads();
  if (n_stats > 0) {
    for (int i = threadIdx.x; i < p.N; i += V2_THREADS) {
      atomicAdd(p.col_sum + i, s_stats[i]);
      atomicAdd(p.col_sumsq + i, s_stats[n_stats + i]);
    }
  }
  if (warp == 1) tmem_dealloc(tmem_base, tmem_cols);
  if (n_stats > 0 && p.has_tail) bn_tail_run(p.tail, p.col_sum, p.col_sumsq, p.N, gridDim.x, tmem_slot + 1);   // spare word next to the TMEM slot
}

// ------------------------------------------------------------------------------------------------ host side
static bool gemm_v2_enabled() {
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("RT1B200_GEMM_PERSISTENT");
    on = (e == nullptr || e[0] != '0') ? 1 : 0;
  }
  return on == 1;
}

static long v2_min_tiles() {
  static long v = -1;
  if (v < 0) {
    const char* e = getenv("RT1B200_GEMM_PERSISTENT_MIN_TILES");
    v = e ? atol(e) : 1;
  }
  return v;
}

// bf16-output GEMMs: persistent kernel, one CTA per SM, TMA-store epilogue
static int launch_gemm_persistent(GemmParams& p, const void* A, long long lda, const void* B, long long ldb, void* out,
                                  long long ldo, const void* bias, const void* residual, long long ldr, float alpha,
                                  float dropout_p, unsigned seed, float* col_sum, float* col_sumsq, cudaStream_t stream,
                                  const Rt1BnTail* tail = nullptr) {
  const int M = p.M, N = p.N, K = p.K;
  p.has_tail = tail != nullptr ? 1 : 0;
  if (tail != nullptr) p.tail = *tail;
  const int bn = N <= 128 ? ((N + 15) / 16) * 16 : 128;
  p.block_n = bn;
  p.kb_per_split = (K + BLOCK_K - 1) / BLOCK_K;
  p.out = out; p.ldo = ldo; p.out_fp32 = 0; p.atomic_out = 0;
  p.bias = (const bf16*)bias; p.residual = (const bf16*)residual; p.ldr = ldr;
  p.alpha = alpha; p.dropout_p = dropout_p; p.seed = seed; p.seed_epoch = dropout_p > 0.f ? rt1_seed_epoch_ptr() : nullptr;
  p.col_sum = col_sum; p.col_sumsq = col_sumsq;
  {
    const char* e = getenv("RT1B200_GEMM_DBG");
    p.dbg = e ? atoi(e) : 0;
  }
  const int n_tiles_h = (N + bn - 1) / bn, b_bytes_h = b_tile_bytes(bn, p.b_mn);
  const long bres = (long)n_tiles_h * p.kb_per_split * b_bytes_h;
  p.b_resident = (bres <= V2_BRES_MAX && !(p.dbg & 32)) ? 1 : 0;
  const int bres_bytes = p.b_resident ? (int)bres : 0;
  const int stage_bytes = A_TILE_BYTES + (p.b_resident ? 0 : b_bytes_h);
  const int n_stats = col_sum != nullptr ? ((N + 3) & ~3) : 0;
  int stages = p.b_resident ? 6 : 4;
  while (stages > 2 && v2_smem_layout(stages, stage_bytes, bres_bytes, n_stats).total + 1024 > 227 * 1024) --stages;
  p.stages = stages;
  const size_t smem = (size_t)v2_smem_layout(stages, stage_bytes, bres_bytes, n_stats).total + 1024;
  RT1_REQUIRE(smem <= 227 * 1024, "rt1_gemm_bf16: persistent kernel needs %zu bytes of shared memory", smem);

  CUtensorMap ta, tb, to;
  int rc;
  if (!p.a_mn) rc = rt1_make_tmap_bf16_2d(&ta, A, (uint64_t)M, (uint64_t)K, (uint64_t)lda, BLOCK_M, BLOCK_K);
  else rc = rt1_make_tmap_bf16_2d(&ta, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda, BLOCK_K, 64);
  if (rc) return rc;
  if (!p.b_mn) rc = rt1_make_tmap_bf16_2d(&tb, B, (uint64_t)N, (uint64_t)K, (uint64_t)ldb, (uint32_t)bn, BLOCK_K);
  else rc = rt1_make_tmap_bf16_2d(&tb, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb, BLOCK_K, 64);
  if (rc) return rc;
  rc = rt1_make_tmap_bf16_2d(&to, out, (uint64_t)M, (uint64_t)N, (uint64_t)ldo, BLOCK_M, 64);
  if (rc) return rc;

  if ((rc = rt1_ensure_dyn_smem((const void*)gemm_tcgen05_persistent_kernel<true>, smem))) return rc;
  if ((rc = rt1_ensure_dyn_smem((const void*)gemm_tcgen05_persistent_kernel<false>, smem))) return rc;
  const int sm_count = rt1_sm_count();
  const long tiles = (long)((M + BLOCK_M - 1) / BLOCK_M) * ((N + bn - 1) / bn);
  const int grid = (int)(tiles < sm_count ? tiles : sm_count);
  const bool plain = alpha == 1.f && bias == nullptr && residual == nullptr && dropout_p <= 0.f;
  if (plain) rt1_launch(gemm_tcgen05_persistent_kernel<true>, grid, V2_THREADS, smem, stream, ta, tb, to, p);
  else rt1_launch(gemm_tcgen05_persistent_kernel<false>, grid, V2_THREADS, smem, stream, ta, tb, to, p);
  return rt1_check_launch("rt1_gemm_bf16(persistent)");
}

// D[M,N] = alpha * op(A) · op(B)^T (+bias) → dropout → (+residual)
//   a_mn == 0: A is [M, K] row-major (lda = row stride);  a_mn == 1: A is stored as [K, M] row-major
//   b_mn == 0: B is [N, K] row-major (ldb);               b_mn == 1: B is stored as [K, N] row-major
//   out: bf16 (out_fp32 = 0) or fp32; split_k > 1 requires out_fp32 and pre-zeroed output (atomic accumulate)
RT1_API int rt1_gemm_bf16(const void* A, long long lda, int a_mn, const void* B, long long ldb, int b_mn, void* out,
                          long long ldo, int out_fp32, int M, int N, int K, const void* bias, const void* residual,
                          long long ldr, float alpha, float dropout_p, unsigned seed, int split_k, int accumulate,
                          float* col_sum, float* col_sumsq, cudaStream_t stream) {
  RT1_REQUIRE(M > 0 && N > 0 && K > 0, "rt1_gemm_bf16: empty problem %dx%dx%d", M, N, K);
  RT1_REQUIRE(split_k >= 1, "rt1_gemm_bf16: split_k must be >= 1");
  RT1_REQUIRE(split_k == 1 || out_fp32, "rt1_gemm_bf16: split-K needs an fp32 output");
  RT1_REQUIRE(!(accumulate && !out_fp32), "rt1_gemm_bf16: accumulate needs an fp32 output");
  // every split (and every accumulating call) runs the whole epilogue on its partial sum: an epilogue term would be
  // applied `splits` times and column statistics would be taken over partial sums
  RT1_REQUIRE(!((split_k > 1 || accumulate) && (bias != nullptr || residual != nullptr || dropout_p > 0.f || col_sum != nullptr ||
                                                col_sumsq != nullptr)),
              "rt1_gemm_bf16: split-K / accumulate cannot be combined with bias, residual, dropout or column statistics");
  GemmParams p;
  p.M = M; p.N = N; p.K = K; p.a_mn = a_mn; p.b_mn = b_mn;
  // The persistent kernel pays ~10 us of set-up (TMEM allocation, barrier init, pipeline fill) and runs ONE CTA per SM:
  // it wins when every SM gets a run of tiles (the huge-M 1x1 convolutions), the tile-per-CTA kernel below — 8 CTAs
  // per SM, 64-column tiles — wins for the decoder and the late, small-M layers.
  const long tiles128 = (long)((M + 127) / 128) * ((N + 127) / 128);
  if (!out_fp32 && split_k == 1 && !accumulate && ((uintptr_t)out & 15) == 0 && (ldo * 2) % 16 == 0 && tiles128 >= v2_min_tiles() &&
      gemm_v2_enabled())
    return launch_gemm_persistent(p, A, lda, B, ldb, out, ldo, bias, residual, ldr, alpha, dropout_p, seed, col_sum, col_sumsq, stream);
  // tile width: whole N when it fits in one UMMA (<= 256), else 128 — rounded up to the UMMA N granule of 16
  // Tile width: at most 64 accumulator columns per CTA.  These GEMMs are short-K and memory-bound; what hides
  // the TMA → MMA → epilogue latency of a tile is the number of CTAs resident per SM, and TMEM allows
  // 512 / (allocated columns) of them — 64 columns → 8 CTAs/SM (a 144-wide tile would allocate 256 → 2 CTAs/SM,
  // measured 9x slower).  A is re-read N/64 times, from L2.
  int bn;
  if (N <= 64) bn = ((N + 15) / 16) * 16;
  else {
    const int parts = (N + 63) / 64;
    bn = (((N + parts - 1) / parts) + 15) / 16 * 16;
  }
  p.block_n = bn;
  const int total_kb = (K + BLOCK_K - 1) / BLOCK_K;
  p.kb_per_split = (total_kb + split_k - 1) / split_k;
  const int splits = (total_kb + p.kb_per_split - 1) / p.kb_per_split;
  const int stage_bytes = A_TILE_BYTES + b_tile_bytes(bn, b_mn);
  int stages = p.kb_per_split < 4 ? p.kb_per_split : 4;
  if (stages < 1) stages = 1;
  // the column-statistics epilogue reuses the operand stages as a 128 x 33 fp32 transpose buffer
  while (col_sum != nullptr && stages * stage_bytes < (128 * 17 + 256) * 4) ++stages;
  p.stages = stages;
  p.out = out; p.ldo = ldo; p.out_fp32 = out_fp32; p.atomic_out = (splits > 1 || accumulate) ? 1 : 0;
  p.bias = (const bf16*)bias; p.residual = (const bf16*)residual; p.ldr = ldr;
  p.alpha = alpha; p.dropout_p = dropout_p; p.seed = seed; p.seed_epoch = dropout_p > 0.f ? rt1_seed_epoch_ptr() : nullptr;
  p.col_sum = col_sum; p.col_sumsq = col_sumsq; p.dbg = 0; p.b_resident = 0; p.has_tail = 0;

  CUtensorMap ta, tb;
  int rc;
  if (!a_mn) rc = rt1_make_tmap_bf16_2d(&ta, A, (uint64_t)M, (uint64_t)K, (uint64_t)lda, BLOCK_M, BLOCK_K);
  else rc = rt1_make_tmap_bf16_2d(&ta, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda, BLOCK_K, 64);
  if (rc) return rc;
  if (!b_mn) rc = rt1_make_tmap_bf16_2d(&tb, B, (uint64_t)N, (uint64_t)K, (uint64_t)ldb, (uint32_t)bn, BLOCK_K);
  else rc = rt1_make_tmap_bf16_2d(&tb, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb, BLOCK_K, 64);
  if (rc) return rc;

  const size_t smem = (size_t)stages * stage_bytes + (2 * stages + 2) * 8 + 1024;
  if ((rc = rt1_ensure_dyn_smem((const void*)gemm_tcgen05_kernel, smem))) return rc;
  dim3 grid(((M + BLOCK_M - 1) / BLOCK_M) * ((N + bn - 1) / bn), 1, splits);
  rt1_launch(gemm_tcgen05_kernel, grid, 128, smem, stream, ta, tb, p);
  return rt1_check_launch("rt1_gemm_bf16");
}

// bf16-output GEMM whose epilogue takes the BatchNorm statistics of the result AND finalises them (see bn_tail.cuh): every
// 1x1 convolution of the image tokenizer in training mode.  Shapes the persistent kernel does not take (unaligned output)
// fall back to the plain call followed by the finalize launch.
RT1_API int rt1_gemm_bf16_bn(const void* A, long long lda, int a_mn, const void* B, long long ldb, int b_mn, void* out, long long ldo,
                             int M, int N, int K, float* col_sum, float* col_sumsq, const Rt1BnTail* tail, cudaStream_t stream) {
  RT1_REQUIRE(M > 0 && N > 0 && K > 0, "rt1_gemm_bf16_bn: empty problem %dx%dx%d", M, N, K);
  RT1_REQUIRE(col_sum != nullptr && col_sumsq != nullptr && tail != nullptr && tail->ticket != nullptr,
              "rt1_gemm_bf16_bn: statistics buffers, tail and ticket are required");
  const long tiles128 = (long)((M + 127) / 128) * ((N + 127) / 128);
  if (((uintptr_t)out & 15) == 0 && (ldo * 2) % 16 == 0 && tiles128 >= v2_min_tiles() && gemm_v2_enabled()) {
    GemmParams p;
    p.M = M; p.N = N; p.K = K; p.a_mn = a_mn; p.b_mn = b_mn;
    return launch_gemm_persistent(p, A, lda, B, ldb, out, ldo, nullptr, nullptr, 0, 1.f, 0.f, 0u, col_sum, col_sumsq, stream, tail);
  }
  int rc = rt1_gemm_bf16(A, lda, a_mn, B, ldb, b_mn, out, ldo, 0, M, N, K, nullptr, nullptr, 0, 1.f, 0.f, 0u, 1, 0, col_sum, col_sumsq,
                         stream);
  if (rc) return rc;
  return rt1_bn_finalize(col_sum, col_sumsq, tail->count, tail->gamma, tail->beta, tail->running_mean, tail->running_var,
                         tail->momentum, tail->eps, 1, tail->a, tail->b, tail->mean, tail->rstd, N, stream);
}

// column sums of a bf16 [M, N] matrix → bf16 [N] (bias gradients).  32 columns x 32 row lanes per CTA, four loads in flight
// per thread: the first version walked M / 8 rows per thread one dependent 2-byte load at a time (12.5 us for 528 x 512).
constexpr int COLSUM_LANES = 32;
__global__ void __launch_bounds__(32 * COLSUM_LANES) colsum_kernel(const bf16* __restrict__ x, long ld, int M, int N, bf16* __restrict__ out) {
  pdl_enter();
  __shared__ float sm[COLSUM_LANES][33];
  const int cl = threadIdx.x & 31, col = blockIdx.x * 32 + cl, r0 = threadIdx.x >> 5;
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
  if (col < N) {
    const bf16* px = x + col;
    int r = r0;
    for (; r + 3 * COLSUM_LANES < M; r += 4 * COLSUM_LANES) {
      const bf16 v0 = px[(long)r * ld], v1 = px[(long)(r + COLSUM_LANES) * ld], v2 = px[(long)(r + 2 * COLSUM_LANES) * ld],
                 v3 = px[(long)(r + 3 * COLSUM_LANES) * ld];
      s0 += __bfloat162float(v0);
      s1 += __bfloat162float(v1);
      s2 += __bfloat162float(v2);
      s3 += __bfloat162float(v3);
    }
    for (; r < M; r += COLSUM_LANES) s0 += __bfloat162float(px[(long)r * ld]);
  }
  sm[r0][cl] = (s0 + s1) + (s2 + s3);
  __syncthreads();
  if (r0 == 0 && col < N) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < COLSUM_LANES; ++i) t += sm[i][cl];
    out[col] = __float2bfloat16_rn(t);
  }
}
RT1_API int rt1_colsum_bf16(const void* x, long long ld, int M, int N, void* out, cudaStream_t stream) {
  rt1_launch(colsum_kernel, rt1_cdiv(N, 32), 32 * COLSUM_LANES, 0, stream, (const bf16*)x, ld, M, N, (bf16*)out);
  return rt1_check_launch("rt1_colsum_bf16");
}

// fp32 → bf16 cast of a small buffer (split-K weight-gradient workspace → gradient slot)
__global__ void cast_f32_bf16_kernel(const float* __restrict__ in, bf16* __restrict__ out, long n) {
  pdl_enter();
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = __float2bfloat16_rn(in[i]);
}
RT1_API int rt1_cast_f32_bf16(const float* in, void* out, long long n, cudaStream_t stream) {
  rt1_launch(cast_f32_bf16_kernel, rt1_cdiv(n, 256), 256, 0, stream, in, (bf16*)out, n);
  return rt1_check_launch("rt1_cast_f32_bf16");
}
