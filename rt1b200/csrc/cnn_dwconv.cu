// Depthwise k x k convolution (k in {3,5}, stride in {1,2}, padding (k-1)/2) in NHWC bf16: forward, data gradient and
// weight gradient as ONE tile kernel family (replaces cuDNN's depthwise fprop/dgrad/wgrad behind
// film_efficientnet_encoder.py:198-208 of the reference).
//
// Design (sm_100a):
//   * A CTA walks a contiguous run of work items (channel chunk of <= 64 channels, frame, 16-wide spatial tile).  The
//     item's source tile INCLUDING its halo is fetched by one 4-D TMA box load (cp.async.bulk.tensor.4d) into shared
//     memory; coordinates outside the image arrive as zeros, which is the convolution's padding — no bounds tests in the
//     inner loop.  A ring of 2-4 stages: the loads of the next items are in flight while item i is computed; items are
//     walked with incremental cursors (no division per item), tile pointers keep their shared-memory provenance (LDS/STS).
//   * lane = channel pair, warp = position inside the tile.  A thread owns a PX x PY block of destination pixels for two
//     channels; it reads every source row of its footprint ONCE (one conflict-free 4-byte LDS per pixel: the 32 lanes of a
//     warp read 32 consecutive words) and uses it for all PY x K row/tap combinations, so a source element costs one
//     shared-memory read per thread instead of K reads through L1 (the round-1 kernels were L1/TEX-bound at 4-7 % of
//     DRAM bandwidth).  All tap indices are compile-time after unrolling.
//   * the two channels of a thread are the two halves of a packed fp32 pair: every multiply-add is one FFMA2
//     (fma.rn.f32x2), the only way to reach the full fp32 rate on sm_100 (a 3-register FFMA issues every other cycle).
//   * fusions: BatchNorm+SiLU of the producer can be applied to the tile in shared memory (each source element once per
//     CTA, packed pairs, zero padding re-imposed afterwards); the forward emits the per-channel sum / sum of squares the next
//     BatchNorm needs and, through its last CTA, that BatchNorm's scale / shift (bn_tail.cuh); the data gradient multiplies by SiLU'(BN(x)) and emits the BatchNorm-backward sums; the weight
//     gradient keeps its K*K accumulators in registers across all items of a chunk.
#include "cnn_common.cuh"
#include "sm100.cuh"
#include "bn_tail.cuh"

using namespace sm100;

namespace {

typedef unsigned long long f2;  // two fp32 values (channels c, c+1) in one 64-bit register pair

__device__ __forceinline__ f2 f2_pack(float lo, float hi) {
  f2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void f2_unpack(f2 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ f2 f2_fma(f2 a, f2 b, f2 c) {
  f2 d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ f2 f2_mul(f2 a, f2 b) {
  f2 d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ f2 f2_add(f2 a, f2 b) {
  f2 d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
// bf16x2 word (low half = even channel) → packed fp32 pair: two ALU ops, no conversion instruction
__device__ __forceinline__ f2 bf2_to_f2(uint32_t u) { return f2_pack(__uint_as_float(u << 16), __uint_as_float(u & 0xffff0000u)); }
__device__ __forceinline__ uint32_t f2_to_bf2(float lo, float hi) {
  const bf162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<const uint32_t*>(&h);
}

enum { DW_FWD = 0, DW_BWD = 1, DW_WGRAD = 2 };

// Geometry of one instantiation.  "Destination" pixels are what a thread holds in registers (forward: outputs; data
// gradient: input gradients; weight gradient: the output gradients it multiplies with), "source" is the haloed tile.
template <int K, int S, int MODE>
struct DwGeom {
  static constexpr int PAD = (K - 1) / 2;
  static constexpr int QLO = PAD / 2;  // data gradient, stride 2: source rows below destination row / 2
  static constexpr int PX = 4;
  static constexpr int PY = (S == 2 && MODE != DW_BWD) ? 2 : 4;
  static constexpr int POSX = 4, POSY = 2;  // 8 warps
  static constexpr int TX = PX * POSX, TY = PY * POSY;
  static constexpr int QX = MODE == DW_BWD ? (S == 1 ? PX + K - 1 : PX / 2 + PAD) : (PX - 1) * S + K;
  static constexpr int QY = MODE == DW_BWD ? (S == 1 ? PY + K - 1 : PY / 2 + PAD) : (PY - 1) * S + K;
  static constexpr int STEPX = MODE == DW_BWD ? (S == 1 ? PX : PX / 2) : PX * S;
  static constexpr int STEPY = MODE == DW_BWD ? (S == 1 ? PY : PY / 2) : PY * S;
  static constexpr int SX = (POSX - 1) * STEPX + QX, SY = (POSY - 1) * STEPY + QY;
  // kernel tap that links destination offset p and source offset q of a thread (valid iff in [0, K))
  __host__ __device__ static constexpr int tap(int p, int q) {
    return MODE == DW_BWD ? (S == 1 ? p + K - 1 - q : p + PAD + 2 * QLO - 2 * q) : q - S * p;
  }
  // source coordinate of the tile origin for destination tile origin d0 (d0 is a multiple of TX / TY, hence even)
  __host__ __device__ static constexpr int src0(int d0) { return MODE == DW_BWD ? (S == 1 ? d0 - PAD : d0 / 2 - QLO) : d0 * S - PAD; }
};

struct DwParams {
  const bf16* w;           // [C][K][K]
  bf16* dst;               // forward: y [N,DH,DW,C]; data gradient: din [N,DH,DW,C]
  const float* a_in;       // forward / weight gradient: BN scale of the producer, applied with SiLU to the source tile (or null)
  const float* b_in;       // data gradient: BN scale/shift of the tensor whose SiLU' multiplies the result (or null)
  const float* mean;       // data gradient with BN
  const float* rstd;
  float* s1;               // forward: sum; data gradient: sum dh            (fp32 [C], caller-zeroed, or null)
  float* s2;               // forward: sum of squares; data gradient: sum dh*xhat
  float* dw;               // weight gradient: fp32 [K*K][C], zeroed by the host wrapper
  int N, C, SH, SW, DH, DW;  // source / destination spatial sizes
  int tiles_x, tiles_y, cw, cww;
  int items, per_cta;
  int src_bytes, aux_bytes, src_region, aux_region;
  int stages;              // depth of the tile ring (2 .. DW_MAX_STAGES)
  int has_tail;            // forward: the CTA that lands the last statistics finalises the BatchNorm (bn_tail.cuh)
  Rt1BnTail tail;
};

constexpr int DW_THREADS = 256;
constexpr int DW_MAX_STAGES = 4;

template <int K, int S, int MODE, int CWW>
__global__ void __launch_bounds__(DW_THREADS, (K == 3 && MODE != DW_BWD) ? 3 : 2)
dw_tile_kernel(const __grid_constant__ CUtensorMap src_map, const __grid_constant__ CUtensorMap aux_map, const DwParams p) {
  using G = DwGeom<K, S, MODE>;
  constexpr int PX = G::PX, PY = G::PY, QX = G::QX, QY = G::QY, SX = G::SX, TX = G::TX, TY = G::TY;
  extern __shared__ __align__(128) uint8_t dw_dyn[];
  __shared__ __align__(8) uint64_t full_bar[DW_MAX_STAGES];
  __shared__ __align__(16) float red[8 * 32 * 2 * (MODE == DW_WGRAD ? K : 2)];   // read back as float4

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int posx = warp % G::POSX, posy = warp / G::POSX;
  const int cww = CWW ? CWW : p.cww;
  // 128-byte alignment by POINTER arithmetic: rounding through an integer hides from the compiler that this is shared
  // memory, and every tile access becomes a generic LD/ST with 64-bit address arithmetic (ncu: IADD3 + IMAD = 20 % of the
  // executed instructions) instead of LDS/STS with immediate offsets
  uint8_t* const base = dw_dyn + ((128u - (smem_u32(dw_dyn) & 127u)) & 127u);
  const int stage_bytes = p.src_region + p.aux_region;

  const int it0 = blockIdx.x * p.per_cta, it1 = min(p.items, it0 + p.per_cta);
  if (it0 >= it1) return;
  if (tid == 0) {
    for (int s = 0; s < p.stages; ++s) mbar_init(smem_u32(&full_bar[s]), 1);
    fence_barrier_init();
    tma_prefetch_desc(&src_map);
    if (p.aux_bytes) tma_prefetch_desc(&aux_map);
  }
  __syncthreads();
  pdl_enter();

  // Item = (chunk, frame, tile row, tile column), column fastest.  Two cursors walk the CTA's run of items by increments —
  // one for the item being computed, one for the item being prefetched — instead of three divisions and three remainders
  // per item and thread (~150 of the ~950 instructions a warp spent per 3x3 item).
  struct Cursor {
    int tx, ty, n, chunk;
    __device__ __forceinline__ void init(int it, const DwParams& q) {
      tx = it % q.tiles_x;
      int r = it / q.tiles_x;
      ty = r % q.tiles_y;
      r /= q.tiles_y;
      n = r % q.N;
      chunk = r / q.N;
    }
    __device__ __forceinline__ void next(const DwParams& q) {
      if (++tx == q.tiles_x) {
        tx = 0;
        if (++ty == q.tiles_y) {
          ty = 0;
          if (++n == q.N) {
            n = 0;
            ++chunk;
          }
        }
      }
    }
  };
  auto issue = [&](const Cursor& c, int stage) {
    const uint32_t bar = smem_u32(&full_bar[stage]);
    const uint32_t dst = smem_u32(base + (size_t)stage * stage_bytes);
    fence_proxy_async();
    mbar_expect_tx(bar, (uint32_t)(p.src_bytes + p.aux_bytes));
    tma_load_4d(dst, &src_map, bar, c.chunk * p.cw, G::src0(c.tx * TX), G::src0(c.ty * TY), c.n);
    if (p.aux_bytes) tma_load_4d(dst + p.src_region, &aux_map, bar, c.chunk * p.cw, c.tx * TX, c.ty * TY, c.n);
  };
  // ring of p.stages tiles: items k+1 .. k+stages-1 are in flight while item k is computed
  Cursor cur, pre;
  cur.init(it0, p);
  pre = cur;                         // only thread 0 advances and uses `pre`
  if (tid == 0)
    for (int s = 0; s < p.stages - 1 && it0 + s < it1; ++s) {
      issue(pre, s);
      pre.next(p);
    }

  // per-chunk thread state
  f2 w[K][K];                       // forward / data gradient: taps of this lane's channel pair; weight gradient: accumulators
  f2 pa = 0, pb = 0, pmu = 0, prs = 0;   // BN constants of the lane's channel pair
  f2 st1 = 0, st2 = 0;              // running statistics of the chunk
  int cur_chunk = -1, c_lo = 0;     // c_lo = first channel of this lane
  bool chan_ok = false;

  auto flush = [&]() {
    // combine the 8 warps (= positions) that share a channel pair, then one atomic per channel (and tap)
    if (MODE == DW_WGRAD) {
#pragma unroll
      for (int ky = 0; ky < K; ++ky) {
        __syncthreads();
#pragma unroll
        for (int kx = 0; kx < K; ++kx) {
          float lo, hi;
          f2_unpack(w[ky][kx], lo, hi);
          red[((warp * K + kx) * 32 + lane) * 2 + 0] = lo;
          red[((warp * K + kx) * 32 + lane) * 2 + 1] = hi;
          w[ky][kx] = 0;
        }
        __syncthreads();
        for (int i = tid; i < K * 64; i += DW_THREADS) {
          const int kx = i >> 6, l2 = i & 63;
          float s = 0.f;
#pragma unroll
          for (int wp = 0; wp < 8; ++wp) s += red[((wp * K + kx) * 32) * 2 + l2];
          const int ch = cur_chunk * p.cw + l2;
          if (l2 < p.cw && ch < p.C) atomicAdd(p.dw + (size_t)(ky * K + kx) * p.C + ch, s);
        }
      }
      __syncthreads();
    } else if (p.s1 != nullptr) {
      __syncthreads();
      float a0, a1, b0, b1;
      f2_unpack(st1, a0, a1);
      f2_unpack(st2, b0, b1);
      float4* r4 = reinterpret_cast<float4*>(red);
      r4[warp * 32 + lane] = make_float4(a0, a1, b0, b1);
      __syncthreads();
      if (tid < 128) {
        const int l = tid >> 2, comp = tid & 3;
        float s = 0.f;
#pragma unroll
        for (int wp = 0; wp < 8; ++wp) s += red[(wp * 32 + l) * 4 + comp];
        const int cl = 2 * l + (comp & 1), ch = cur_chunk * p.cw + cl;
        if (cl < p.cw && ch < p.C) atomicAdd((comp < 2 ? p.s1 : p.s2) + ch, s);
      }
      __syncthreads();
      st1 = 0;
      st2 = 0;
    }
  };

  int stage = -1, use = 0;            // ring slot of the current item and how many times it has been used before
  for (int it = it0; it < it1; ++it) {
    if (++stage == p.stages) {
      stage = 0;
      ++use;
    }
    __syncthreads();  // every thread is done with the stage of item k-1: it may be refilled (with item k+stages-1)
    if (tid == 0 && it + p.stages - 1 < it1) {
      int ps = stage + p.stages - 1;
      if (ps >= p.stages) ps -= p.stages;
      issue(pre, ps);
      pre.next(p);
    }
    const int tx = cur.tx, ty = cur.ty, n = cur.n, chunk = cur.chunk;
    cur.next(p);
    if (chunk != cur_chunk) {
      if (cur_chunk >= 0) flush();
      cur_chunk = chunk;
      c_lo = chunk * p.cw + 2 * lane;
      chan_ok = 2 * lane < p.cw && c_lo < p.C;
      if (MODE == DW_WGRAD) {
#pragma unroll
        for (int ky = 0; ky < K; ++ky)
#pragma unroll
          for (int kx = 0; kx < K; ++kx) w[ky][kx] = 0;
      } else {
        const bf16* wp = p.w + (size_t)c_lo * K * K;
#pragma unroll
        for (int ky = 0; ky < K; ++ky)
#pragma unroll
          for (int kx = 0; kx < K; ++kx)
            w[ky][kx] = chan_ok ? f2_pack(__bfloat162float(wp[ky * K + kx]), __bfloat162float(wp[K * K + ky * K + kx])) : (f2)0;
      }
      if (p.a_in != nullptr) {
        pa = chan_ok ? f2_pack(p.a_in[c_lo], p.a_in[c_lo + 1]) : (f2)0;
        pb = chan_ok ? f2_pack(p.b_in[c_lo], p.b_in[c_lo + 1]) : (f2)0;
        if (MODE == DW_BWD) {
          pmu = chan_ok ? f2_pack(p.mean[c_lo], p.mean[c_lo + 1]) : (f2)0;
          prs = chan_ok ? f2_pack(p.rstd[c_lo], p.rstd[c_lo + 1]) : (f2)0;
        }
      }
    }
    uint32_t* const tile = reinterpret_cast<uint32_t*>(base + (size_t)stage * stage_bytes);
    const uint32_t* const aux = reinterpret_cast<const uint32_t*>(base + (size_t)stage * stage_bytes + p.src_region);
    mbar_wait(smem_u32(&full_bar[stage]), (uint32_t)(use & 1));

    if (MODE != DW_BWD && p.a_in != nullptr) {
      // SiLU(BN(.)) of the producer applied to the tile in place, once per element; padding stays zero
      const int sy0 = G::src0(ty * TY), sx0 = G::src0(tx * TX);
      // SiLU(u) = u sigmoid(u) = h tanh(h) + h with h = u / 2 = (a/2) x + b/2: one packed FMA, two MUFU.TANH, one packed FMA
      // per channel pair (the scalar form was 2 FFMA + 1 FMUL + 1 MUFU per element, 11 % of the kernel's instructions)
      const f2 half = f2_pack(0.5f, 0.5f);
      const f2 ah = f2_mul(pa, half), bh = f2_mul(pb, half);
      if (lane < cww) {
        static_assert(SX >= 8, "the pixel walk below wraps at most once per step of 8");
        int ry = 0, rx = warp;                    // pixel `pix` = (ry, rx), advanced by 8 without a division
        uint32_t* q = tile + warp * cww + lane;
        for (int pix = warp; pix < G::SY * SX; pix += 8, q += 8 * cww) {
          const bool inside = (unsigned)(sy0 + ry) < (unsigned)p.SH && (unsigned)(sx0 + rx) < (unsigned)p.SW;
          rx += 8;
          if (rx >= SX) {
            rx -= SX;
            ++ry;
          }
          const f2 h = f2_fma(ah, bf2_to_f2(*q), bh);
          float h0, h1;
          f2_unpack(h, h0, h1);
          const f2 y = f2_fma(h, f2_pack(tanh_approx_f(h0), tanh_approx_f(h1)), h);
          float y0, y1;
          f2_unpack(y, y0, y1);
          *q = inside ? f2_to_bf2(y0, y1) : 0u;
        }
      }
      fence_proxy_async();   // these generic-proxy writes precede the TMA refill of this stage `stages` items later
      __syncthreads();
    }

    const uint32_t* const src = tile + ((posy * G::STEPY) * SX + posx * G::STEPX) * cww + lane;
    const int dx0 = tx * TX + posx * PX, dy0 = ty * TY + posy * PY;

    if (MODE == DW_WGRAD) {
      f2 g[PY][PX];
#pragma unroll
      for (int py = 0; py < PY; ++py)
#pragma unroll
        for (int px = 0; px < PX; ++px) g[py][px] = bf2_to_f2(aux[((posy * PY + py) * TX + posx * PX + px) * cww + lane]);
#pragma unroll
      for (int q = 0; q < QY; ++q) {
        f2 row[QX];
#pragma unroll
        for (int j = 0; j < QX; ++j) row[j] = bf2_to_f2(src[(q * SX + j) * cww]);
#pragma unroll
        for (int py = 0; py < PY; ++py) {
          const int ky = G::tap(py, q);
          if (ky >= 0 && ky < K) {
#pragma unroll
            for (int px = 0; px < PX; ++px)
#pragma unroll
              for (int j = 0; j < QX; ++j) {
                const int kx = G::tap(px, j);
                if (kx >= 0 && kx < K) w[ky][kx] = f2_fma(g[py][px], row[j], w[ky][kx]);
              }
          }
        }
      }
    } else {
      f2 acc[PY][PX];
#pragma unroll
      for (int py = 0; py < PY; ++py)
#pragma unroll
        for (int px = 0; px < PX; ++px) acc[py][px] = 0;
#pragma unroll
      for (int q = 0; q < QY; ++q) {
        f2 row[QX];
#pragma unroll
        for (int j = 0; j < QX; ++j) row[j] = bf2_to_f2(src[(q * SX + j) * cww]);
#pragma unroll
        for (int py = 0; py < PY; ++py) {
          const int ky = G::tap(py, q);
          if (ky >= 0 && ky < K) {
#pragma unroll
            for (int px = 0; px < PX; ++px)
#pragma unroll
              for (int j = 0; j < QX; ++j) {
                const int kx = G::tap(px, j);
                if (kx >= 0 && kx < K) acc[py][px] = f2_fma(w[ky][kx], row[j], acc[py][px]);
              }
          }
        }
      }
      // epilogue: the warp shares its position, so the bounds tests are warp-uniform
      bf16* const out = p.dst + (size_t)n * p.DH * p.DW * p.C + c_lo;
      const bool bwd_bn = MODE == DW_BWD && p.a_in != nullptr;
      const f2 c_half = f2_pack(0.5f, 0.5f), c_one = f2_pack(1.f, 1.f), c_neg1 = f2_pack(-1.f, -1.f);
      const f2 nmr = f2_mul(f2_mul(pmu, prs), c_neg1);      // xhat = x rstd - mean rstd
#pragma unroll
      for (int py = 0; py < PY; ++py) {
        const int dy = dy0 + py;
        if (dy >= p.DH) continue;
#pragma unroll
        for (int px = 0; px < PX; ++px) {
          const int dx = dx0 + px;
          if (dx >= p.DW) continue;
          float lo, hi;
          if (bwd_bn) {
            // dh = dAct * SiLU'(u), u = a x + b; SiLU'(u) = s + u s (1 - s), s = sigmoid(u) = tanh(u/2)/2 + 1/2.  All of it
            // on channel PAIRS (fma.rn.f32x2): 9 packed operations + 2 MUFU per pair instead of ~22 scalar ones — the
            // epilogue was a fifth of the 3x3 data-gradient kernel's instructions.  BatchNorm-backward sums over dh, dh*xhat.
            const f2 x = bf2_to_f2(aux[((posy * PY + py) * TX + posx * PX + px) * cww + lane]);
            const f2 u = f2_fma(pa, x, pb);
            float h0, h1;
            f2_unpack(f2_mul(u, c_half), h0, h1);
            const f2 sg = f2_fma(f2_pack(tanh_approx_f(h0), tanh_approx_f(h1)), c_half, c_half);
            const f2 d = f2_fma(f2_mul(u, sg), f2_fma(sg, c_neg1, c_one), sg);
            const f2 dh = f2_mul(acc[py][px], d);
            st1 = f2_add(st1, dh);
            st2 = f2_fma(dh, f2_fma(x, prs, nmr), st2);
            f2_unpack(dh, lo, hi);
          } else {
            f2_unpack(acc[py][px], lo, hi);
          }
          if (!bwd_bn && MODE == DW_FWD) {
            st1 = f2_add(st1, acc[py][px]);
            st2 = f2_fma(acc[py][px], acc[py][px], st2);
          }
          if (chan_ok) *reinterpret_cast<uint32_t*>(out + ((size_t)dy * p.DW + dx) * p.C) = f2_to_bf2(lo, hi);
        }
      }
    }
  }
  flush();
  if (MODE == DW_FWD && p.has_tail) {   // every CTA of the grid has items
    __shared__ unsigned s_last;
    bn_tail_run(p.tail, p.s1, p.s2, p.C, gridDim.x, &s_last);
  }
}

// scratch [K*K][C] fp32 → weight gradient [C][K*K] bf16
__global__ void dw_wgrad_cast_kernel(const float* __restrict__ scratch, bf16* __restrict__ dw, int C, int KK) {
  pdl_enter();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= C * KK) return;
  const int c = i / KK, t = i % KK;
  dw[i] = __float2bfloat16_rn(scratch[(size_t)t * C + c]);
}

template <int K, int S, int MODE, int CWW>
int launch_dw_inst(const CUtensorMap& src_map, const CUtensorMap& aux_map, const DwParams& p, size_t smem, int grid, cudaStream_t stream) {
  if (int rc = rt1_ensure_dyn_smem((const void*)dw_tile_kernel<K, S, MODE, CWW>, 200 * 1024)) return rc;
  rt1_launch(dw_tile_kernel<K, S, MODE, CWW>, grid, DW_THREADS, smem, stream, src_map, aux_map, p);
  return 0;
}

// src: the haloed operand [N,SH,SW,C]; aux: the operand read at destination coordinates [N,DH,DW,C] (or null)
template <int K, int S, int MODE>
int launch_dw(const void* src, const void* aux, DwParams p, cudaStream_t stream, const char* what) {
  using G = DwGeom<K, S, MODE>;
  RT1_REQUIRE(p.C % 8 == 0, "%s: C=%d must be a multiple of 8", what, p.C);
  p.cw = p.C >= 64 ? 64 : p.C;
  p.cww = p.cw / 2;
  const int chunks = (p.C + p.cw - 1) / p.cw;
  p.tiles_x = (p.DW + G::TX - 1) / G::TX;
  p.tiles_y = (p.DH + G::TY - 1) / G::TY;
  const long items = (long)chunks * p.N * p.tiles_x * p.tiles_y;
  RT1_REQUIRE(items > 0 && items < (1L << 30), "%s: bad problem size", what);
  p.items = (int)items;
  p.src_bytes = G::SY * G::SX * p.cw * 2;
  p.aux_bytes = aux != nullptr ? G::TY * G::TX * p.cw * 2 : 0;
  p.src_region = (p.src_bytes + 127) & ~127;
  p.aux_region = (p.aux_bytes + 127) & ~127;
  const size_t stage_b = (size_t)(p.src_region + p.aux_region);
  size_t smem = 2 * stage_b + 256;  // alignment slack + over-read slack of idle lanes
  RT1_REQUIRE(smem <= 200 * 1024, "%s: tile needs %zu bytes of shared memory", what, smem);
  const size_t static_smem = (MODE == DW_WGRAD ? 8 * 32 * 2 * K : 8 * 32 * 4) * sizeof(float) + 64 + 1024;  // + per-CTA reserve
  int per_sm = (int)((227 * 1024) / (smem + static_smem));
  const int cap = (K == 3 && MODE != DW_BWD) ? 3 : 2;
  per_sm = per_sm < 1 ? 1 : (per_sm > cap ? cap : per_sm);
  // deepen the ring as far as the CTA's share of shared memory allows without costing a resident CTA
  p.stages = 2;
  {
    static int max_stages = -1;
    if (max_stages < 0) {
      const char* e = getenv("RT1B200_DW_STAGES");
      max_stages = e ? atoi(e) : DW_MAX_STAGES;
      if (max_stages < 2) max_stages = 2;
      if (max_stages > DW_MAX_STAGES) max_stages = DW_MAX_STAGES;
    }
    const size_t budget = (227 * 1024) / per_sm - static_smem;
    while (p.stages < max_stages && (size_t)(p.stages + 1) * stage_b + 256 <= budget && (size_t)(p.stages + 1) * stage_b + 256 <= 200 * 1024)
      ++p.stages;
    smem = (size_t)p.stages * stage_b + 256;
  }
  int grid = rt1_sm_count() * per_sm;
  if ((long)grid > items) grid = (int)items;
  p.per_cta = (int)((items + grid - 1) / grid);
  grid = (int)((items + p.per_cta - 1) / p.per_cta);
  CUtensorMap src_map, aux_map;
  int rc = rt1_make_tmap_bf16_nhwc(&src_map, src, (uint64_t)p.N, (uint64_t)p.SH, (uint64_t)p.SW, (uint64_t)p.C, G::SY, G::SX, (uint32_t)p.cw);
  if (rc) return rc;
  if (aux != nullptr) {
    rc = rt1_make_tmap_bf16_nhwc(&aux_map, aux, (uint64_t)p.N, (uint64_t)p.DH, (uint64_t)p.DW, (uint64_t)p.C, G::TY, G::TX, (uint32_t)p.cw);
    if (rc) return rc;
  } else {
    aux_map = src_map;
  }
  if (p.cww == 32) rc = launch_dw_inst<K, S, MODE, 32>(src_map, aux_map, p, smem, grid, stream);
  else rc = launch_dw_inst<K, S, MODE, 0>(src_map, aux_map, p, smem, grid, stream);
  if (rc) return rc;
  return rt1_check_launch(what);
}

#define DW_DISPATCH(MODE, ...)                                                                       \
  do {                                                                                               \
    if (k == 3 && stride == 1) return launch_dw<3, 1, MODE>(__VA_ARGS__);                             \
    if (k == 3 && stride == 2) return launch_dw<3, 2, MODE>(__VA_ARGS__);                             \
    if (k == 5 && stride == 1) return launch_dw<5, 1, MODE>(__VA_ARGS__);                             \
    if (k == 5 && stride == 2) return launch_dw<5, 2, MODE>(__VA_ARGS__);                             \
    RT1_REQUIRE(false, "depthwise conv: unsupported kernel %d stride %d", k, stride);                 \
  } while (0)

int dw_wgrad_impl(const void* dout, const void* x, const float* a_in, const float* b_in, float* scratch, int N, int H, int W, int C,
                  int k, int stride, cudaStream_t stream) {
  const int pad = (k - 1) / 2, Ho = (H + 2 * pad - k) / stride + 1, Wo = (W + 2 * pad - k) / stride + 1;
  DwParams p = {};
  p.a_in = a_in; p.b_in = b_in; p.dw = scratch;
  p.N = N; p.C = C; p.SH = H; p.SW = W; p.DH = Ho; p.DW = Wo;
  DW_DISPATCH(DW_WGRAD, x, dout, p, stream, "rt1_dwconv_wgrad");
}

}  // namespace

// x [N,H,W,C] raw (a_in/b_in = producer BN scale/shift, SiLU applied on load) or already-activated (a_in = null);
// y [N,Ho,Wo,C] raw output; sum/sumsq fp32 [C] (caller-zeroed) or null.
RT1_API int rt1_dwconv_fwd(const void* x, const float* a_in, const float* b_in, const void* w, void* y, float* sum, float* sumsq,
                           int N, int H, int W, int C, int k, int stride, cudaStream_t stream) {
  const int pad = (k - 1) / 2, Ho = (H + 2 * pad - k) / stride + 1, Wo = (W + 2 * pad - k) / stride + 1;
  DwParams p = {};
  p.w = (const bf16*)w; p.dst = (bf16*)y; p.a_in = a_in; p.b_in = b_in; p.s1 = sum; p.s2 = sumsq;
  p.N = N; p.C = C; p.SH = H; p.SW = W; p.DH = Ho; p.DW = Wo;
  DW_DISPATCH(DW_FWD, x, nullptr, p, stream, "rt1_dwconv_fwd");
}

// rt1_dwconv_fwd whose last-arriving CTA also finalises the BatchNorm of the output (scale/shift, mean/rstd, running stats)
RT1_API int rt1_dwconv_fwd_bn(const void* x, const float* a_in, const float* b_in, const void* w, void* y, float* sum, float* sumsq,
                              const Rt1BnTail* tail, int N, int H, int W, int C, int k, int stride, cudaStream_t stream) {
  RT1_REQUIRE(sum != nullptr && sumsq != nullptr && tail != nullptr && tail->ticket != nullptr,
              "rt1_dwconv_fwd_bn: statistics buffers, tail and ticket are required");
  const int pad = (k - 1) / 2, Ho = (H + 2 * pad - k) / stride + 1, Wo = (W + 2 * pad - k) / stride + 1;
  DwParams p = {};
  p.w = (const bf16*)w; p.dst = (bf16*)y; p.a_in = a_in; p.b_in = b_in; p.s1 = sum; p.s2 = sumsq;
  p.N = N; p.C = C; p.SH = H; p.SW = W; p.DH = Ho; p.DW = Wo;
  p.has_tail = 1; p.tail = *tail;
  DW_DISPATCH(DW_FWD, x, nullptr, p, stream, "rt1_dwconv_fwd_bn");
}

// dout [N,Ho,Wo,C]; x_raw [N,H,W,C] + a_in/b_in/mean/rstd (or all null when the conv input was not BN+SiLU);
// din [N,H,W,C]: d/d(x_raw-normalised pre-activation) "dh" when BN is present, else d/d(input);
// sums fp32 [2,C] (caller-zeroed) receives sum dh and sum dh*xhat (BN case).
RT1_API int rt1_dwconv_bwd_data(const void* dout, const void* w, const void* x_raw, const float* a_in, const float* b_in,
                                const float* mean, const float* rstd, void* din, float* sums, int N, int H, int W, int C, int k,
                                int stride, cudaStream_t stream) {
  const int pad = (k - 1) / 2, Ho = (H + 2 * pad - k) / stride + 1, Wo = (W + 2 * pad - k) / stride + 1;
  DwParams p = {};
  p.w = (const bf16*)w; p.dst = (bf16*)din; p.a_in = a_in; p.b_in = b_in; p.mean = mean; p.rstd = rstd;
  p.s1 = (a_in != nullptr) ? sums : nullptr; p.s2 = (a_in != nullptr && sums != nullptr) ? sums + C : nullptr;
  p.N = N; p.C = C; p.SH = Ho; p.SW = Wo; p.DH = H; p.DW = W;
  const void* aux = a_in != nullptr ? x_raw : nullptr;
  DW_DISPATCH(DW_BWD, dout, aux, p, stream, "rt1_dwconv_bwd_data");
}

// scratch: fp32 [k*k*C], zeroed by this call; dw: bf16 [C, k, k]
RT1_API int rt1_dwconv_wgrad(const void* dout, const void* x, const float* a_in, const float* b_in, float* scratch, void* dw, int N,
                             int H, int W, int C, int k, int stride, cudaStream_t stream) {
  RT1_CHECK_CUDA(cudaMemsetAsync(scratch, 0, sizeof(float) * (size_t)k * k * C, stream));
  int rc = dw_wgrad_impl(dout, x, a_in, b_in, scratch, N, H, W, C, k, stride, stream);
  if (rc) return rc;
  rt1_launch(dw_wgrad_cast_kernel, rt1_cdiv((long)C * k * k, 256), 256, 0, stream, scratch, (bf16*)dw, C, k * k);
  return rt1_check_launch("rt1_dwconv_wgrad(cast)");
}

// ---- stem as a tensor-core GEMM: im2col of the 3x3 stride-2 window -------------------------------------------------
// x [N,H,W,8] bf16 (channels 0..2 used) → cols [N*Ho*Wo, 32] bf16, column ci*9 + ky*3 + kx (the flattening of the torch
// weight [CO,3,3,3]); columns 27..31 are zero.  64 bytes per output pixel of extra traffic buy a K = 32 tcgen05 GEMM for the
// forward (with the BatchNorm statistics in its epilogue) and a split-K GEMM for the weight gradient; a direct CUDA-core stem
// (round 1) ran at ~7 % of DRAM bandwidth.
__global__ void __launch_bounds__(256) stem_im2col_kernel(const bf16* __restrict__ x, bf16* __restrict__ cols, int H, int W, int Ho,
                                                          int Wo, long total) {
  pdl_enter();
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int ox = (int)(i % Wo), oy = (int)((i / Wo) % Ho);
  const long n = i / ((long)Wo * Ho);
  const bf16* xin = x + n * (long)H * W * 8;
  __align__(16) bf16 v[32];
#pragma unroll
  for (int t = 0; t < 32; ++t) v[t] = __float2bfloat16_rn(0.f);
#pragma unroll
  for (int ky = 0; ky < 3; ++ky) {
    const int iy = oy * 2 - 1 + ky;
#pragma unroll
    for (int kx = 0; kx < 3; ++kx) {
      const int ix = ox * 2 - 1 + kx;
      if (iy < 0 || iy >= H || ix < 0 || ix >= W) continue;
      const uint2 px = *reinterpret_cast<const uint2*>(xin + ((long)iy * W + ix) * 8);   // channels 0..3
      const bf16* c = reinterpret_cast<const bf16*>(&px);
      v[0 * 9 + ky * 3 + kx] = c[0];
      v[1 * 9 + ky * 3 + kx] = c[1];
      v[2 * 9 + ky * 3 + kx] = c[2];
    }
  }
  uint4* dst = reinterpret_cast<uint4*>(cols + i * 32);
#pragma unroll
  for (int q = 0; q < 4; ++q) dst[q] = reinterpret_cast<const uint4*>(v)[q];
}

RT1_API int rt1_stem_im2col(const void* x, void* cols, int N, int H, int W, cudaStream_t stream) {
  const int Ho = (H + 2 - 3) / 2 + 1, Wo = (W + 2 - 3) / 2 + 1;
  const long total = (long)N * Ho * Wo;
  rt1_launch(stem_im2col_kernel, rt1_cdiv(total, 256), 256, 0, stream, (const bf16*)x, (bf16*)cols, H, W, Ho, Wo, total);
  return rt1_check_launch("rt1_stem_im2col");
}
