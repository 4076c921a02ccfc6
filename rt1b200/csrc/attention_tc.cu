// Masked attention of the RT-1 decoder on the 5th-generation tensor cores (replaces the two batched
// matmuls + masked_fill + softmax + dropout of reference transformer.py:82-109).
//
// One CTA (128 threads) per (batch, head).  The whole problem — S <= 128 tokens (66 in RT-1), dk in {64, 128} — lives
// in one CTA: Q/K/V (and dO) tiles are TMA-loaded straight out of the fused projection output qkv [b, S, 3, h, dk]
// into 128-byte-swizzled shared memory, every contraction is a tcgen05.mma with its accumulator in TMEM, and thread t
// owns TMEM lane t = one query row, so mask, softmax and dropout are thread-local (no shuffles, no shared-memory
// score matrix).  P and dS are written back to shared memory as bf16 in the canonical UMMA layout; the SAME bytes serve
// as a K-major operand (rows = queries, K = keys) and as an MN-major operand (K rows = queries, M = keys), so
//   forward : S = Q K^T          → softmax/dropout → O = P V
//   backward: S = Q K^T, dP = dO V^T → P, dS       → dV = P^T dO, dQ = dS K, dK = dS^T Q
// need no transposes: Q, K, V, dO are each loaded once and re-described.  Rows beyond S inside the 128-row tiles hold
// whatever follows in memory (finite) or TMA zero fill; P/dS are written with exact zeros wherever such a row or column
// would enter a contraction.  Dropout uses the same stateless hash as the generic kernel (attention.cu), so masks are
// reproducible on the host and identical in forward and backward.
#include "sm100.cuh"

using namespace sm100;

namespace {

constexpr int TC_THREADS = 128;
constexpr int TILE_ROWS = 128;            // query rows (= TMEM lanes = UMMA M)
constexpr int CHUNK_BYTES = TILE_ROWS * 128;   // one 64-column swizzle chunk of a 128-row tile

struct AttnTcParams {
  int B, S, H, DK, SP;       // SP = S rounded up to 16 (UMMA K / N granule)
  float scale, p_drop;
  uint32_t seed;
  const uint32_t* seed_epoch;
  const uint8_t* mask;       // [S, S] or null
  bf16* out;                 // forward: [B, S, H*DK]
  bf16* dqkv;                // backward: [B, S, 3, H, DK]
};

// 16 fp32 accumulator columns of this thread's TMEM lane
__device__ __forceinline__ void ld16(uint32_t taddr, float* v) {
  uint32_t r[16];
  tmem_ld_32x16(taddr, r);
  tmem_ld_wait();
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// write 8 consecutive bf16 (columns j0 .. j0+7, j0 % 8 == 0) of row r into a [128 x 128] K-major SW128 tile pair
__device__ __forceinline__ void st_tile8(uint8_t* tile, int r, int j0, const float* v) {
  const int chunk = j0 >> 6, unit = (j0 & 63) >> 3;
  uint8_t* p = tile + chunk * CHUNK_BYTES + r * 128 + ((unit ^ (r & 7)) << 4);
  *reinterpret_cast<bf16x8*>(p) = pack8(v);
}

// this thread's mask row as a bit set (bit j = key j visible); no mask = everything visible
__device__ __forceinline__ void load_mask_bits(const uint8_t* mask, int row, int S, uint32_t (&bits)[4]) {
#pragma unroll
  for (int w = 0; w < 4; ++w) bits[w] = 0xffffffffu;
  if (mask == nullptr || row >= S) return;
  const uint8_t* m = mask + (size_t)row * S;
#pragma unroll
  for (int w = 0; w < 4; ++w) {
    uint32_t b = 0;
    for (int j = 0; j < 32; ++j) {
      const int col = w * 32 + j;
      if (col < S && m[col] != 0) b |= 1u << j;
    }
    bits[w] = b;
  }
}
__device__ __forceinline__ bool bit_of(const uint32_t (&bits)[4], int j) { return (bits[j >> 5] >> (j & 31)) & 1u; }

// row maximum and 1 / sum(exp) of the masked, scaled scores of this thread's row (TMEM columns [0, SP) at `ts`)
__device__ __forceinline__ void row_softmax_stats(uint32_t ts, int S, int SP, float scale, const uint32_t (&bits)[4], float& mx,
                                                  float& inv) {
  mx = -INFINITY;
  for (int c = 0; c < SP; c += 16) {
    float v[16];
    ld16(ts + c, v);
#pragma unroll
    for (int i = 0; i < 16; ++i)
      if (c + i < S) mx = fmaxf(mx, bit_of(bits, c + i) ? v[i] * scale : -1e9f);
  }
  float sum = 0.f;
  for (int c = 0; c < SP; c += 16) {
    float v[16];
    ld16(ts + c, v);
#pragma unroll
    for (int i = 0; i < 16; ++i)
      if (c + i < S) sum += __expf((bit_of(bits, c + i) ? v[i] * scale : -1e9f) - mx);
  }
  inv = 1.f / sum;
}

struct Smem {
  uint8_t *q, *k, *v, *go, *p, *ds;
  uint64_t* bars;
  uint32_t* tmem_slot;
};
__device__ __forceinline__ Smem carve(uint8_t* raw, int DK, int SP, bool bwd) {
  uint8_t* base = raw + ((1024u - (smem_u32(raw) & 1023u)) & 1023u);   // pointer arithmetic keeps the shared address space
  const int nch = DK / 64;
  Smem s;
  s.q = base;
  s.k = s.q + nch * CHUNK_BYTES;
  s.v = s.k + nch * SP * 128;
  uint8_t* nxt = s.v + nch * SP * 128;
  nxt += (1024u - (smem_u32(nxt) & 1023u)) & 1023u;
  s.go = nxt;
  if (bwd) nxt += nch * CHUNK_BYTES;
  s.p = nxt;
  nxt += 2 * CHUNK_BYTES;
  s.ds = nxt;
  if (bwd) nxt += 2 * CHUNK_BYTES;
  s.bars = reinterpret_cast<uint64_t*>(nxt);
  s.tmem_slot = reinterpret_cast<uint32_t*>(s.bars + 4);
  return s;
}
size_t tc_smem_bytes(int DK, int SP, bool bwd) {
  const size_t nch = DK / 64;
  size_t b = nch * CHUNK_BYTES + 2 * nch * SP * 128 + 1024;
  if (bwd) b += nch * CHUNK_BYTES;
  b += 2 * CHUNK_BYTES * (bwd ? 2 : 1);
  return b + 64 + 1024;
}

// D[tmem] = A · B^T over K = 16 * ksteps, both operands K-major tiles of 128-byte rows (64-column chunks `a_chunk` / `b_chunk` apart)
__device__ __forceinline__ void mma_kmajor(uint32_t d, uint32_t a, int a_chunk, uint32_t b, int b_chunk, int ksteps, uint32_t idesc) {
  for (int k = 0; k < ksteps; ++k) {
    const uint64_t da = umma_smem_desc(a + (k >> 2) * a_chunk + (k & 3) * 32, 16, 1024);
    const uint64_t db = umma_smem_desc(b + (k >> 2) * b_chunk + (k & 3) * 32, 16, 1024);
    umma_bf16(d, da, db, idesc, k != 0);
  }
}

__global__ void __launch_bounds__(TC_THREADS) attn_tc_fwd_kernel(const __grid_constant__ CUtensorMap map_q,
                                                                 const __grid_constant__ CUtensorMap map_kv, const AttnTcParams p) {
  extern __shared__ __align__(1024) uint8_t attn_raw[];
  const Smem sm = carve(attn_raw, p.DK, p.SP, false);
  const int tid = threadIdx.x, warp = tid >> 5;
  const int head = blockIdx.x, b = blockIdx.y;
  const int nch = p.DK / 64, S = p.S, SP = p.SP;
  const uint32_t bar_load = smem_u32(sm.bars), bar_s = smem_u32(sm.bars + 1), bar_o = smem_u32(sm.bars + 2);

  if (tid == 0) {
    tma_prefetch_desc(&map_q);
    tma_prefetch_desc(&map_kv);
    mbar_init(bar_load, 1);
    mbar_init(bar_s, 1);
    mbar_init(bar_o, 1);
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(smem_u32(sm.tmem_slot), 256);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *sm.tmem_slot;
  pdl_enter();
  const uint32_t t_s = tmem, t_o = tmem + 128;

  if (tid == 0) {
    mbar_expect_tx(bar_load, (uint32_t)(nch * (CHUNK_BYTES + 2 * SP * 128)));
    const int row0 = b * S;
    for (int c = 0; c < nch; ++c) {
      tma_load_2d(smem_u32(sm.q + c * CHUNK_BYTES), &map_q, bar_load, (0 * p.H + head) * p.DK + 64 * c, row0);
      tma_load_2d(smem_u32(sm.k + c * SP * 128), &map_kv, bar_load, (1 * p.H + head) * p.DK + 64 * c, row0);
      tma_load_2d(smem_u32(sm.v + c * SP * 128), &map_kv, bar_load, (2 * p.H + head) * p.DK + 64 * c, row0);
    }
    mbar_wait(bar_load, 0);
    tc_fence_after();
    mma_kmajor(t_s, smem_u32(sm.q), CHUNK_BYTES, smem_u32(sm.k), SP * 128, p.DK / 16, umma_idesc_bf16(128, SP, 0, 0));
    umma_commit(bar_s);
  }
  uint32_t bits[4];
  load_mask_bits(p.mask, tid, S, bits);
  __syncwarp();
  mbar_wait(bar_s, 0);
  tc_fence_after();

  // tcgen05.ld is warp-collective (.sync.aligned): every thread runs the loops, rows >= S just do not store
  const uint32_t lane_addr = (uint32_t)(warp * 32) << 16;
  const bool row_ok = tid < S;
  {
    float mx, inv;
    row_softmax_stats(t_s + lane_addr, S, SP, p.scale, bits, mx, inv);
    const uint32_t thresh = rt1_drop_threshold(p.p_drop);
    const float ks = p.p_drop > 0.f ? 1.f / (1.f - p.p_drop) : 1.f;
    const uint32_t seed = p.seed + (p.seed_epoch != nullptr ? __ldg(p.seed_epoch) : 0u);
    const uint32_t rng_base = (uint32_t)((b * p.H + head) * S * S + tid * S);
    for (int c = 0; c < SP; c += 16) {
      float v[16];
      ld16(t_s + lane_addr + c, v);
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const int j = c + i;
        float pr = 0.f;
        if (row_ok && j < S) {
          pr = __expf((bit_of(bits, j) ? v[i] * p.scale : -1e9f) - mx) * inv;
          if (p.p_drop > 0.f) pr = rt1_hash_u32(rng_base + (uint32_t)j, seed) >= thresh ? pr * ks : 0.f;
        }
        v[i] = pr;
      }
      st_tile8(sm.p, tid, c, v);          // rows >= S get exact zeros: their output rows are never stored
      st_tile8(sm.p, tid, c + 8, v + 8);
    }
  }
  fence_proxy_async();   // P was written through the generic proxy, the tensor core reads it through the async proxy
  tc_fence_before();
  __syncthreads();
  if (tid == 0) {
    tc_fence_after();
    const uint32_t idesc = umma_idesc_bf16(128, p.DK, 0, 1);   // A = P (K-major over keys), B = V (MN-major: rows are keys)
    for (int k = 0; k < SP / 16; ++k) {
      const uint64_t da = umma_smem_desc(smem_u32(sm.p) + (k >> 2) * CHUNK_BYTES + (k & 3) * 32, 16, 1024);
      const uint64_t db = umma_smem_desc(smem_u32(sm.v) + k * 2048, SP * 128, 1024);
      umma_bf16(t_o, da, db, idesc, k != 0);
    }
    umma_commit(bar_o);
  }
  __syncwarp();
  mbar_wait(bar_o, 0);
  tc_fence_after();
  {
    bf16* o = p.out + ((size_t)(b * S + (row_ok ? tid : 0)) * p.H + head) * p.DK;
    for (int c = 0; c < p.DK; c += 16) {
      float v[16];
      ld16(t_o + lane_addr + c, v);
      if (row_ok) {
        *reinterpret_cast<bf16x8*>(o + c) = pack8(v);
        *reinterpret_cast<bf16x8*>(o + c + 8) = pack8(v + 8);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 256);
}

__global__ void __launch_bounds__(TC_THREADS) attn_tc_bwd_kernel(const __grid_constant__ CUtensorMap map_q,
                                                                 const __grid_constant__ CUtensorMap map_kv,
                                                                 const __grid_constant__ CUtensorMap map_go, const AttnTcParams p) {
  extern __shared__ __align__(1024) uint8_t attn_raw[];
  const Smem sm = carve(attn_raw, p.DK, p.SP, true);
  const int tid = threadIdx.x, warp = tid >> 5;
  const int head = blockIdx.x, b = blockIdx.y;
  const int nch = p.DK / 64, S = p.S, SP = p.SP, DK = p.DK;
  const uint32_t bar_load = smem_u32(sm.bars), bar_1 = smem_u32(sm.bars + 1), bar_2 = smem_u32(sm.bars + 2);

  if (tid == 0) {
    tma_prefetch_desc(&map_q);
    tma_prefetch_desc(&map_kv);
    tma_prefetch_desc(&map_go);
    mbar_init(bar_load, 1);
    mbar_init(bar_1, 1);
    mbar_init(bar_2, 1);
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(smem_u32(sm.tmem_slot), 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *sm.tmem_slot;
  pdl_enter();
  const uint32_t t_s = tmem, t_dp = tmem + 128, t_dv = tmem + 256, t_dq = tmem + 384, t_dk = tmem;   // dK reuses the S columns

  if (tid == 0) {
    mbar_expect_tx(bar_load, (uint32_t)(nch * (2 * CHUNK_BYTES + 2 * SP * 128)));
    const int row0 = b * S;
    for (int c = 0; c < nch; ++c) {
      tma_load_2d(smem_u32(sm.q + c * CHUNK_BYTES), &map_q, bar_load, (0 * p.H + head) * DK + 64 * c, row0);
      tma_load_2d(smem_u32(sm.k + c * SP * 128), &map_kv, bar_load, (1 * p.H + head) * DK + 64 * c, row0);
      tma_load_2d(smem_u32(sm.v + c * SP * 128), &map_kv, bar_load, (2 * p.H + head) * DK + 64 * c, row0);
      tma_load_2d(smem_u32(sm.go + c * CHUNK_BYTES), &map_go, bar_load, head * DK + 64 * c, row0);
    }
    mbar_wait(bar_load, 0);
    tc_fence_after();
    const uint32_t idesc = umma_idesc_bf16(128, SP, 0, 0);
    mma_kmajor(t_s, smem_u32(sm.q), CHUNK_BYTES, smem_u32(sm.k), SP * 128, DK / 16, idesc);      // S  = Q K^T
    mma_kmajor(t_dp, smem_u32(sm.go), CHUNK_BYTES, smem_u32(sm.v), SP * 128, DK / 16, idesc);    // dP = dO V^T (pre-dropout-mask)
    umma_commit(bar_1);
  }
  uint32_t bits[4];
  load_mask_bits(p.mask, tid, S, bits);
  __syncwarp();
  mbar_wait(bar_1, 0);
  tc_fence_after();

  // tcgen05.ld is warp-collective (.sync.aligned): every thread runs the loops; rows >= S write exact zeros — query
  // rows S .. SP-1 enter the contractions over queries (dV, dK), the rest is never read
  const uint32_t lane_addr = (uint32_t)(warp * 32) << 16;
  const bool row_ok = tid < S;
  {
    float mx, inv;
    row_softmax_stats(t_s + lane_addr, S, SP, p.scale, bits, mx, inv);
    const uint32_t thresh = rt1_drop_threshold(p.p_drop);
    const float ks = p.p_drop > 0.f ? 1.f / (1.f - p.p_drop) : 1.f;
    const uint32_t seed = p.seed + (p.seed_epoch != nullptr ? __ldg(p.seed_epoch) : 0u);
    const uint32_t rng_base = (uint32_t)((b * p.H + head) * S * S + tid * S);
    // rowdot = sum_j dP_j P_j with dP_j = keep_j * ks * <dO_i, V_j>
    float rowdot = 0.f;
    for (int c = 0; c < SP; c += 16) {
      float v[16], g[16];
      ld16(t_s + lane_addr + c, v);
      ld16(t_dp + lane_addr + c, g);
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const int j = c + i;
        if (row_ok && j < S) {
          const float pr = __expf((bit_of(bits, j) ? v[i] * p.scale : -1e9f) - mx) * inv;
          const bool keep = p.p_drop <= 0.f || rt1_hash_u32(rng_base + (uint32_t)j, seed) >= thresh;
          rowdot += keep ? g[i] * ks * pr : 0.f;
        }
      }
    }
    for (int c = 0; c < SP; c += 16) {
      float v[16], g[16];
      ld16(t_s + lane_addr + c, v);
      ld16(t_dp + lane_addr + c, g);
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const int j = c + i;
        float pd = 0.f, dsv = 0.f;
        if (row_ok && j < S) {
          const float pr = __expf((bit_of(bits, j) ? v[i] * p.scale : -1e9f) - mx) * inv;
          const bool keep = p.p_drop <= 0.f || rt1_hash_u32(rng_base + (uint32_t)j, seed) >= thresh;
          pd = keep ? pr * ks : 0.f;
          dsv = pr * ((keep ? g[i] * ks : 0.f) - rowdot) * p.scale;
        }
        v[i] = pd;
        g[i] = dsv;
      }
      st_tile8(sm.p, tid, c, v);
      st_tile8(sm.p, tid, c + 8, v + 8);
      st_tile8(sm.ds, tid, c, g);
      st_tile8(sm.ds, tid, c + 8, g + 8);
    }
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  if (tid == 0) {
    tc_fence_after();
    const int ks16 = SP / 16;
    // dV = P^T dO : A = P read MN-major (M = keys, K = queries), B = dO read MN-major (N = dk, K = queries)
    {
      const uint32_t idesc = umma_idesc_bf16(128, DK, 1, 1);
      for (int k = 0; k < ks16; ++k)
        umma_bf16(t_dv, umma_smem_desc(smem_u32(sm.p) + k * 2048, CHUNK_BYTES, 1024),
                  umma_smem_desc(smem_u32(sm.go) + k * 2048, CHUNK_BYTES, 1024), idesc, k != 0);
    }
    // dQ = dS K : A = dS K-major (K = keys), B = K read MN-major (N = dk, K rows = keys)
    {
      const uint32_t idesc = umma_idesc_bf16(128, DK, 0, 1);
      for (int k = 0; k < ks16; ++k)
        umma_bf16(t_dq, umma_smem_desc(smem_u32(sm.ds) + (k >> 2) * CHUNK_BYTES + (k & 3) * 32, 16, 1024),
                  umma_smem_desc(smem_u32(sm.k) + k * 2048, SP * 128, 1024), idesc, k != 0);
    }
    // dK = dS^T Q : A = dS read MN-major (M = keys, K = queries), B = Q read MN-major
    {
      const uint32_t idesc = umma_idesc_bf16(128, DK, 1, 1);
      for (int k = 0; k < ks16; ++k)
        umma_bf16(t_dk, umma_smem_desc(smem_u32(sm.ds) + k * 2048, CHUNK_BYTES, 1024),
                  umma_smem_desc(smem_u32(sm.q) + k * 2048, CHUNK_BYTES, 1024), idesc, k != 0);
    }
    umma_commit(bar_2);
  }
  __syncwarp();
  mbar_wait(bar_2, 0);
  tc_fence_after();
  {
    const size_t tok = (size_t)(b * S + (row_ok ? tid : 0)) * 3 * p.H * DK + (size_t)head * DK;
    bf16* dq = p.dqkv + tok;
    bf16* dk = dq + (size_t)p.H * DK;
    bf16* dv = dq + (size_t)2 * p.H * DK;
    for (int c = 0; c < DK; c += 16) {
      float v[16];
      ld16(t_dq + lane_addr + c, v);
      if (row_ok) {
        *reinterpret_cast<bf16x8*>(dq + c) = pack8(v);
        *reinterpret_cast<bf16x8*>(dq + c + 8) = pack8(v + 8);
      }
      ld16(t_dk + lane_addr + c, v);
      if (row_ok) {
        *reinterpret_cast<bf16x8*>(dk + c) = pack8(v);
        *reinterpret_cast<bf16x8*>(dk + c + 8) = pack8(v + 8);
      }
      ld16(t_dv + lane_addr + c, v);
      if (row_ok) {
        *reinterpret_cast<bf16x8*>(dv + c) = pack8(v);
        *reinterpret_cast<bf16x8*>(dv + c + 8) = pack8(v + 8);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

}  // namespace

bool rt1_attention_tc_supported(int S, int DK, const void* a, const void* b) {
  static int off = -1;
  if (off < 0) {
    const char* e = getenv("RT1B200_ATTN_TC");
    off = (e != nullptr && e[0] == '0') ? 1 : 0;
  }
  return off == 0 && (DK == 64 || DK == 128) && S >= 1 && S <= 128 && ((uintptr_t)a & 15) == 0 && ((uintptr_t)b & 15) == 0;
}

int rt1_attention_tc_fwd(const void* qkv, const unsigned char* mask, void* out, int B, int S, int H, int DK, float p_drop,
                         unsigned seed, cudaStream_t stream) {
  AttnTcParams p = {};
  p.B = B; p.S = S; p.H = H; p.DK = DK; p.SP = (S + 15) / 16 * 16;
  p.scale = 1.f / sqrtf((float)DK); p.p_drop = p_drop; p.seed = seed;
  p.seed_epoch = p_drop > 0.f ? rt1_seed_epoch_ptr() : nullptr;
  p.mask = mask; p.out = (bf16*)out;
  CUtensorMap mq, mkv;
  const uint64_t rows = (uint64_t)B * S, cols = (uint64_t)3 * H * DK;
  int rc = rt1_make_tmap_bf16_2d(&mq, qkv, rows, cols, cols, TILE_ROWS, 64);
  if (rc) return rc;
  rc = rt1_make_tmap_bf16_2d(&mkv, qkv, rows, cols, cols, (uint32_t)p.SP, 64);
  if (rc) return rc;
  const size_t smem = tc_smem_bytes(DK, p.SP, false);
  if ((rc = rt1_ensure_dyn_smem((const void*)attn_tc_fwd_kernel, smem))) return rc;
  rt1_launch(attn_tc_fwd_kernel, dim3(H, B), TC_THREADS, smem, stream, mq, mkv, p);
  return rt1_check_launch("rt1_attention_fwd(tcgen05)");
}

int rt1_attention_tc_bwd(const void* qkv, const unsigned char* mask, const void* dout, void* dqkv, int B, int S, int H, int DK,
                         float p_drop, unsigned seed, cudaStream_t stream) {
  AttnTcParams p = {};
  p.B = B; p.S = S; p.H = H; p.DK = DK; p.SP = (S + 15) / 16 * 16;
  p.scale = 1.f / sqrtf((float)DK); p.p_drop = p_drop; p.seed = seed;
  p.seed_epoch = p_drop > 0.f ? rt1_seed_epoch_ptr() : nullptr;
  p.mask = mask; p.dqkv = (bf16*)dqkv;
  CUtensorMap mq, mkv, mgo;
  const uint64_t rows = (uint64_t)B * S, cols = (uint64_t)3 * H * DK, ocols = (uint64_t)H * DK;
  int rc = rt1_make_tmap_bf16_2d(&mq, qkv, rows, cols, cols, TILE_ROWS, 64);
  if (rc) return rc;
  rc = rt1_make_tmap_bf16_2d(&mkv, qkv, rows, cols, cols, (uint32_t)p.SP, 64);
  if (rc) return rc;
  rc = rt1_make_tmap_bf16_2d(&mgo, dout, rows, ocols, ocols, TILE_ROWS, 64);
  if (rc) return rc;
  const size_t smem = tc_smem_bytes(DK, p.SP, true);
  if ((rc = rt1_ensure_dyn_smem((const void*)attn_tc_bwd_kernel, smem))) return rc;
  rt1_launch(attn_tc_bwd_kernel, dim3(H, B), TC_THREADS, smem, stream, mq, mkv, mgo, p);
  return rt1_check_launch("rt1_attention_bwd(tcgen05)");
}
