// Depthwise k x k convolution, second mapping: TWO channels per thread, the whole k x k filter in registers.
//
// ncu on the 8-channels-per-thread kernels (cnn_dwconv.cu): L1/TEX 89 % busy, DRAM 3 %.  A thread owned 8 channels, so the
// k x k x 8 filter (200 floats at k = 5) could not live in registers and was re-read from shared memory per kernel row and
// item; with no broadcast possible (every thread has other channels) those reads cost as many L1 wavefronts as the input.
// Here a thread owns a channel PAIR (one 4-byte bf16x2 load per tap column):
//   * its 2 x k x k filter values sit in registers for the life of the thread — no weight traffic in the loop at all;
//   * a warp reads 128 contiguous bytes per load instruction (one L1 wavefront instead of four);
//   * PX = 8 output pixels of a row per item: a row of PX + k - 1 loads feeds 8 outputs (1.5 loads per output row-tap set
//     at k = 5 instead of 2);
//   * the weight gradient keeps all k x k taps of its pair in registers, so one CTA covers the whole filter (the
//     8-channel kernel had to split kernel rows over CTAs and re-read both operands k times).
// Only the plain variants used by the training schedule are implemented (input already activated, stride-1 data
// gradient); everything else falls back to cnn_dwconv.cu.
//
// MEASURED (B200, one training step, 1 GPU): correct (same tests) but SLOWER than the 8-channel kernels — forward 5x5
// 1.90 vs 1.47 ms, data gradient 5x5 2.73 vs 1.81 ms, weight gradient 5x5 2.19 vs 1.69 ms per step: four times as many
// load instructions (4 bytes each) move the bottleneck from L1 wavefronts to instruction issue.  Kept as an opt-in
// (RT1B200_DW_PAIR=1) and as the record of the experiment; the 8-channel kernels stay the default.
#include <stdlib.h>

#include "cnn_common.cuh"

#define DW2_THREADS 256
#define DW2_MAXP 128   // channel pairs per CTA chunk

struct PairMap {
  int np;      // channel pairs handled by this CTA
  int cp;      // this thread's channel pair (global index)
  int r0;      // first item offset of this thread
  int lanes;   // item stride
  bool active;
};
__device__ __forceinline__ PairMap make_pair_map(int C2, int chunk_idx) {
  PairMap m;
  const int nchunks = (C2 + DW2_MAXP - 1) / DW2_MAXP;
  int np = (C2 + nchunks - 1) / nchunks;
  const int chunk0 = chunk_idx * np;
  np = min(np, C2 - chunk0);
  m.np = np;
  m.lanes = DW2_THREADS / np;
  m.active = (int)threadIdx.x < m.lanes * np;
  m.cp = chunk0 + threadIdx.x % np;
  m.r0 = threadIdx.x / np;
  return m;
}
static inline int dw2_chunks(int C) { return (C / 2 + DW2_MAXP - 1) / DW2_MAXP; }

__device__ __forceinline__ float2 ld_pair(const bf16* p) { return __bfloat1622float2(*reinterpret_cast<const bf162*>(p)); }
__device__ __forceinline__ void st_pair(bf16* p, float2 v) { *reinterpret_cast<bf162*>(p) = __floats2bfloat162_rn(v.x, v.y); }

// sum a per-thread float2 over the lanes of the CTA that share a channel pair, then one atomic per channel
__device__ __forceinline__ void cta_pair_reduce_atomic(const PairMap& m, float2 acc, float2* sm, float* gdst) {
  __syncthreads();
  sm[threadIdx.x] = m.active ? acc : make_float2(0.f, 0.f);
  __syncthreads();
  if ((int)threadIdx.x < m.np) {
    float2 s = make_float2(0.f, 0.f);
    for (int r = 0; r < m.lanes; ++r) {
      const float2 t = sm[r * m.np + threadIdx.x];
      s.x += t.x;
      s.y += t.y;
    }
    const int chunk0 = m.cp - (threadIdx.x % m.np);
    atomicAdd(gdst + (size_t)(chunk0 + threadIdx.x) * 2, s.x);
    atomicAdd(gdst + (size_t)(chunk0 + threadIdx.x) * 2 + 1, s.y);
  }
}

// FLIP = false: y = conv(x, w) (stride S).  FLIP = true (S == 1): data gradient = conv(dout, w flipped), followed by the
// SiLU'(BN(x_raw)) factor and the BatchNorm-backward sums when x_raw is given.
template <int K, int S, int PX, bool FLIP>
__global__ void __launch_bounds__(DW2_THREADS, 2) dw2_conv_kernel(const bf16* __restrict__ x, const bf16* __restrict__ w,
                                                                  bf16* __restrict__ y, float* __restrict__ sum, float* __restrict__ sumsq,
                                                                  const bf16* __restrict__ x_raw, const float* __restrict__ a_in,
                                                                  const float* __restrict__ b_in, const float* __restrict__ mean,
                                                                  const float* __restrict__ rstd, int H, int W, int Ho, int Wo, int C) {
  __shared__ float2 red[DW2_THREADS];
  const PairMap m = make_pair_map(C / 2, blockIdx.z);
  constexpr int PAD = (K - 1) / 2, NIN = (PX - 1) * S + K;
  const int n = blockIdx.y, XG = (Wo + PX - 1) / PX, items = Ho * XG;
  int ib, ie;
  cta_rows(items, ib, ie);
  float2 s1 = make_float2(0.f, 0.f), s2 = make_float2(0.f, 0.f);
  if (m.active) {
    float2 wf[K * K];
#pragma unroll
    for (int t = 0; t < K * K; ++t) {
      const int ts = FLIP ? (K * K - 1 - t) : t;   // 180-degree rotation of the filter for the data gradient
      wf[t] = make_float2(__bfloat162float(w[(size_t)(m.cp * 2) * K * K + ts]), __bfloat162float(w[(size_t)(m.cp * 2 + 1) * K * K + ts]));
    }
    const bool has_bn = FLIP && x_raw != nullptr;
    float2 fa = make_float2(0.f, 0.f), fb = fa, mu = fa, rs = fa;
    if (has_bn) {
      fa = *reinterpret_cast<const float2*>(a_in + m.cp * 2);
      fb = *reinterpret_cast<const float2*>(b_in + m.cp * 2);
      mu = *reinterpret_cast<const float2*>(mean + m.cp * 2);
      rs = *reinterpret_cast<const float2*>(rstd + m.cp * 2);
    }
    const bf16* xin = x + (size_t)n * H * W * C + (size_t)m.cp * 2;
    const size_t out_off = (size_t)n * Ho * Wo * C + (size_t)m.cp * 2;
    for (int it = ib + m.r0; it < ie; it += m.lanes) {
      const int oy = it / XG, ox0 = (it - oy * XG) * PX;
      float2 acc[PX];
#pragma unroll
      for (int p = 0; p < PX; ++p) acc[p] = make_float2(0.f, 0.f);
      const int ix_base = ox0 * S - PAD;
#pragma unroll
      for (int ky = 0; ky < K; ++ky) {
        const int iy = oy * S - PAD + ky;
        if (iy < 0 || iy >= H) continue;
        const bf16* xrow = xin + (size_t)iy * W * C;
#pragma unroll
        for (int j = 0; j < NIN; ++j) {
          const int ix = ix_base + j;
          if (ix < 0 || ix >= W) continue;
          const float2 v = ld_pair(xrow + (size_t)ix * C);
#pragma unroll
          for (int p = 0; p < PX; ++p) {
            const int kx = j - p * S;
            if (kx >= 0 && kx < K) {
              acc[p].x = fmaf(wf[ky * K + kx].x, v.x, acc[p].x);
              acc[p].y = fmaf(wf[ky * K + kx].y, v.y, acc[p].y);
            }
          }
        }
      }
#pragma unroll
      for (int p = 0; p < PX; ++p) {
        const int ox = ox0 + p;
        if (ox >= Wo) continue;
        const size_t o = out_off + ((size_t)oy * Wo + ox) * C;
        float2 r = acc[p];
        if (FLIP) {
          if (has_bn) {
            const float2 xr = ld_pair(x_raw + o);
            r.x *= silu_grad_f(fmaf(fa.x, xr.x, fb.x));
            r.y *= silu_grad_f(fmaf(fa.y, xr.y, fb.y));
            s1.x += r.x; s1.y += r.y;
            s2.x += r.x * (xr.x - mu.x) * rs.x;
            s2.y += r.y * (xr.y - mu.y) * rs.y;
          }
        } else {
          s1.x += r.x; s1.y += r.y;
          s2.x = fmaf(r.x, r.x, s2.x); s2.y = fmaf(r.y, r.y, s2.y);
        }
        st_pair(y + o, r);
      }
    }
  }
  if (sum != nullptr) {
    cta_pair_reduce_atomic(m, s1, red, sum);
    cta_pair_reduce_atomic(m, s2, red, sumsq);
  }
}

// dw[ky][kx][c] += sum over (n, oy, ox) dOut * x; scratch layout [K*K][C] fp32
template <int K, int S, int PX>
__global__ void __launch_bounds__(DW2_THREADS, 2) dw2_wgrad_kernel(const bf16* __restrict__ dout, const bf16* __restrict__ x,
                                                                   float* __restrict__ dw, int H, int W, int Ho, int Wo, int C) {
  __shared__ float2 red[DW2_THREADS];
  const PairMap m = make_pair_map(C / 2, blockIdx.z);
  constexpr int PAD = (K - 1) / 2, NIN = (PX - 1) * S + K;
  const int n = blockIdx.y, XG = (Wo + PX - 1) / PX, items = Ho * XG;
  int ib, ie;
  cta_rows(items, ib, ie);
  float2 acc[K * K];
#pragma unroll
  for (int t = 0; t < K * K; ++t) acc[t] = make_float2(0.f, 0.f);
  if (m.active) {
    const bf16* xin = x + (size_t)n * H * W * C + (size_t)m.cp * 2;
    const bf16* g = dout + (size_t)n * Ho * Wo * C + (size_t)m.cp * 2;
    for (int it = ib + m.r0; it < ie; it += m.lanes) {
      const int oy = it / XG, ox0 = (it - oy * XG) * PX;
      float2 gd[PX];
#pragma unroll
      for (int p = 0; p < PX; ++p) gd[p] = (ox0 + p < Wo) ? ld_pair(g + ((size_t)oy * Wo + ox0 + p) * C) : make_float2(0.f, 0.f);
      const int ix_base = ox0 * S - PAD;
#pragma unroll
      for (int ky = 0; ky < K; ++ky) {
        const int iy = oy * S - PAD + ky;
        if (iy < 0 || iy >= H) continue;
        const bf16* xrow = xin + (size_t)iy * W * C;
#pragma unroll
        for (int j = 0; j < NIN; ++j) {
          const int ix = ix_base + j;
          if (ix < 0 || ix >= W) continue;
          const float2 v = ld_pair(xrow + (size_t)ix * C);
#pragma unroll
          for (int p = 0; p < PX; ++p) {
            const int kx = j - p * S;
            if (kx >= 0 && kx < K) {
              acc[ky * K + kx].x = fmaf(gd[p].x, v.x, acc[ky * K + kx].x);
              acc[ky * K + kx].y = fmaf(gd[p].y, v.y, acc[ky * K + kx].y);
            }
          }
        }
      }
    }
  }
#pragma unroll
  for (int t = 0; t < K * K; ++t) cta_pair_reduce_atomic(m, acc[t], red, dw + (size_t)t * C);
}

static int dw2_items_ctas(int N, int items) {
  int want = (148 * 6 + N - 1) / N;
  int maxc = (items + 7) / 8;
  if (want > maxc) want = maxc;
  return want < 1 ? 1 : want;
}

bool rt1_dw2_enabled() {
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("RT1B200_DW_PAIR");
    on = (e != nullptr && e[0] == '1') ? 1 : 0;   // opt-in: measured slower than the 8-channel kernels (see header comment)
  }
  return on == 1;
}

#define DW2_DISPATCH(KERNEL, PXS1, PXS2, ...)                                                        \
  do {                                                                                                \
    if (k == 3 && stride == 1) KERNEL<3, 1, PXS1> __VA_ARGS__;                                         \
    else if (k == 3 && stride == 2) KERNEL<3, 2, PXS2> __VA_ARGS__;                                    \
    else if (k == 5 && stride == 1) KERNEL<5, 1, PXS1> __VA_ARGS__;                                    \
    else if (k == 5 && stride == 2) KERNEL<5, 2, PXS2> __VA_ARGS__;                                    \
    else RT1_REQUIRE(false, "depthwise conv: unsupported kernel %d stride %d", k, stride);            \
  } while (0)

int rt1_dw2_fwd(const void* x, const void* w, void* y, float* sum, float* sumsq, int N, int H, int W, int C, int k, int stride,
                cudaStream_t stream) {
  const int pad = (k - 1) / 2, Ho = (H + 2 * pad - k) / stride + 1, Wo = (W + 2 * pad - k) / stride + 1;
  const int px = stride == 1 ? 8 : 4;
  const int items = Ho * ((Wo + px - 1) / px);
  dim3 grid(dw2_items_ctas(N, items), N, dw2_chunks(C));
#define DW2_FWD(K_, S_, PX_) dw2_conv_kernel<K_, S_, PX_, false><<<grid, DW2_THREADS, 0, stream>>>((const bf16*)x, (const bf16*)w, (bf16*)y, sum, sumsq, nullptr, nullptr, nullptr, nullptr, nullptr, H, W, Ho, Wo, C)
  if (k == 3 && stride == 1) DW2_FWD(3, 1, 8);
  else if (k == 3 && stride == 2) DW2_FWD(3, 2, 4);
  else if (k == 5 && stride == 1) DW2_FWD(5, 1, 8);
  else if (k == 5 && stride == 2) DW2_FWD(5, 2, 4);
  else RT1_REQUIRE(false, "depthwise conv: unsupported kernel %d stride %d", k, stride);
#undef DW2_FWD
  return rt1_check_launch("rt1_dwconv_fwd(pair)");
}

// stride 1 only: din = conv(dout, rot180(w)) [* SiLU'(BN(x_raw)), + BN-backward sums]
int rt1_dw2_bwd_data_s1(const void* dout, const void* w, const void* x_raw, const float* a_in, const float* b_in, const float* mean,
                        const float* rstd, void* din, float* sums, int N, int H, int W, int C, int k, cudaStream_t stream) {
  const int items = H * ((W + 7) / 8);
  dim3 grid(dw2_items_ctas(N, items), N, dw2_chunks(C));
  float* s1 = sums, * s2 = sums ? sums + C : nullptr;
#define DW2_BWD(K_) dw2_conv_kernel<K_, 1, 8, true><<<grid, DW2_THREADS, 0, stream>>>((const bf16*)dout, (const bf16*)w, (bf16*)din, s1, s2, (const bf16*)x_raw, a_in, b_in, mean, rstd, H, W, H, W, C)
  if (k == 3) DW2_BWD(3);
  else if (k == 5) DW2_BWD(5);
  else RT1_REQUIRE(false, "depthwise conv: unsupported kernel %d", k);
#undef DW2_BWD
  return rt1_check_launch("rt1_dwconv_bwd_data(pair)");
}

// scratch fp32 [k*k*C] must be zero on entry
int rt1_dw2_wgrad(const void* dout, const void* x, float* scratch, int N, int H, int W, int C, int k, int stride, cudaStream_t stream) {
  const int pad = (k - 1) / 2, Ho = (H + 2 * pad - k) / stride + 1, Wo = (W + 2 * pad - k) / stride + 1;
  const int px = stride == 1 ? 8 : 4;
  const int items = Ho * ((Wo + px - 1) / px);
  dim3 grid(dw2_items_ctas(N, items), N, dw2_chunks(C));
#define DW2_WG(K_, S_, PX_) dw2_wgrad_kernel<K_, S_, PX_><<<grid, DW2_THREADS, 0, stream>>>((const bf16*)dout, (const bf16*)x, scratch, H, W, Ho, Wo, C)
  if (k == 3 && stride == 1) DW2_WG(3, 1, 8);
  else if (k == 3 && stride == 2) DW2_WG(3, 2, 4);
  else if (k == 5 && stride == 1) DW2_WG(5, 1, 8);
  else if (k == 5 && stride == 2) DW2_WG(5, 2, 4);
  else RT1_REQUIRE(false, "depthwise conv: unsupported kernel %d stride %d", k, stride);
#undef DW2_WG
  return rt1_check_launch("rt1_dwconv_wgrad(pair)");
}
