// Round-1 depthwise kernels (row-strip per thread, operands through L1), kept ONLY as the timing baseline of
// tools/bench_dwconv.py and as a second oracle in the tests: the product path is the tile kernel in cnn_dwconv.cu.
// Exported as rt1_dwconv_v1_*.
// Thread mapping is channel-stationary (cnn_common.cuh) with channel chunks of <= 64 vectors so the chunk's
// k*k*C weights sit in shared memory; each thread produces PX = 4 neighbouring pixels of a row for its 8 channels
// and activates every loaded input exactly once (sliding window in registers).
#include "cnn_common.cuh"

#define DW_PX 4
#define DW_MAXV 64  // channel vectors per CTA chunk

__device__ __forceinline__ ChanMap make_dw_map(int C8, int chunk_idx) {
  ChanMap m;
  const int nchunks = (C8 + DW_MAXV - 1) / DW_MAXV;
  int c8 = (C8 + nchunks - 1) / nchunks;
  const int chunk0 = chunk_idx * c8;
  c8 = min(c8, C8 - chunk0);
  m.c8 = c8;
  m.rows_per_pass = CNN_THREADS / c8;
  m.active = (int)threadIdx.x < m.rows_per_pass * c8;
  m.cv = chunk0 + threadIdx.x % c8;
  m.r0 = threadIdx.x / c8;
  return m;
}
static inline int dw_chunks(int C) { return (C / 8 + DW_MAXV - 1) / DW_MAXV; }

// weights [C, K, K] (torch [C,1,K,K]) → smem [K*K][2][c8] float4 for this chunk: plane 0 holds channels 0-3 of every
// channel vector, plane 1 channels 4-7, so that consecutive threads (consecutive channel vectors) read consecutive
// float4 — conflict-free.  (With 8 consecutive floats per thread every LDS.128 was a 2-way bank conflict and, read
// per tap and per pixel, the weights were 90 % of the kernel's L1/shared traffic: ncu showed L1/TEX at 89 % busy.)
template <int K>
__device__ __forceinline__ void load_dw_weights(float* wsm, const bf16* w, const ChanMap& m) {
  const int chunk0 = m.cv - (threadIdx.x % m.c8);
  const int nch = m.c8 * 8;
  for (int i = threadIdx.x; i < nch * K * K; i += blockDim.x) {
    const int c = i / (K * K), t = i % (K * K);
    wsm[((t * 2 + ((c & 7) >> 2)) * m.c8 + (c >> 3)) * 4 + (c & 3)] = __bfloat162float(w[(size_t)(chunk0 * 8 + c) * K * K + t]);
  }
  __syncthreads();
}
// the K taps of kernel row ky for this thread's 8 channels → registers (reused by all PX pixels of the item)
template <int K>
__device__ __forceinline__ void load_dw_row(const float* wsm, int c8, int cvl, int ky, float (&wk)[K][8]) {
  const float4* w4 = reinterpret_cast<const float4*>(wsm);
#pragma unroll
  for (int kx = 0; kx < K; ++kx) {
    const float4 lo = w4[((ky * K + kx) * 2 + 0) * c8 + cvl], hi = w4[((ky * K + kx) * 2 + 1) * c8 + cvl];
    wk[kx][0] = lo.x; wk[kx][1] = lo.y; wk[kx][2] = lo.z; wk[kx][3] = lo.w;
    wk[kx][4] = hi.x; wk[kx][5] = hi.y; wk[kx][6] = hi.z; wk[kx][7] = hi.w;
  }
}

template <int K, int S>
__global__ void __launch_bounds__(CNN_THREADS, 2) dwconv_fwd_kernel(const bf16* __restrict__ x, const float* __restrict__ a_in,
                                                                 const float* __restrict__ b_in, const bf16* __restrict__ w,
                                                                 bf16* __restrict__ y, float* __restrict__ sum, float* __restrict__ sumsq,
                                                                 int H, int W, int Ho, int Wo, int C) {
  extern __shared__ float dyn_sm[];
  float* wsm = dyn_sm;                       // [K*K][c8*8]
  float* red = dyn_sm + K * K * DW_MAXV * 8;  // CNN_THREADS * 8
  const ChanMap m = make_dw_map(C / 8, blockIdx.z);
  load_dw_weights<K>(wsm, w, m);
  constexpr int PAD = (K - 1) / 2, NIN = (DW_PX - 1) * S + K;
  const int n = blockIdx.y, XG = (Wo + DW_PX - 1) / DW_PX, items = Ho * XG;
  int ib, ie;
  cta_rows(items, ib, ie);
  float s1[8] = {0, 0, 0, 0, 0, 0, 0, 0}, s2[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (m.active) {
    const int cvl = threadIdx.x % m.c8;
    float fa[8], fb[8];
    const bool has_bn = a_in != nullptr;
    if (has_bn) {
      load8f(a_in + m.cv * 8, fa);
      load8f(b_in + m.cv * 8, fb);
    }
    const bf16* xin = x + (size_t)n * H * W * C + (size_t)m.cv * 8;
    bf16* yout = y + (size_t)n * Ho * Wo * C + (size_t)m.cv * 8;
    for (int it = ib + m.r0; it < ie; it += m.rows_per_pass) {
      const int oy = it / XG, ox0 = (it % XG) * DW_PX;
      float acc[DW_PX][8];
#pragma unroll
      for (int p = 0; p < DW_PX; ++p)
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[p][k] = 0.f;
#pragma unroll 1
      for (int ky = 0; ky < K; ++ky) {
        const int iy = oy * S - PAD + ky;
        if (iy < 0 || iy >= H) continue;
        const int ix_base = ox0 * S - PAD;
        float wk[K][8];
        load_dw_row<K>(wsm, m.c8, cvl, ky, wk);
#pragma unroll
        for (int j = 0; j < NIN; ++j) {
          const int ix = ix_base + j;
          if (ix < 0 || ix >= W) continue;
          float v[8];
          load8bf(xin + ((size_t)iy * W + ix) * C, v);
          if (has_bn) {
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = silu_f(fmaf(fa[k], v[k], fb[k]));
          }
#pragma unroll
          for (int p = 0; p < DW_PX; ++p) {
            const int kx = j - p * S;
            if (kx >= 0 && kx < K) {
#pragma unroll
              for (int k = 0; k < 8; ++k) acc[p][k] = fmaf(wk[kx][k], v[k], acc[p][k]);
            }
          }
        }
      }
#pragma unroll
      for (int p = 0; p < DW_PX; ++p) {
        const int ox = ox0 + p;
        if (ox < Wo) {
          store8bf(yout + ((size_t)oy * Wo + ox) * C, acc[p]);
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            s1[k] += acc[p][k];
            s2[k] += acc[p][k] * acc[p][k];
          }
        }
      }
    }
  }
  if (sum != nullptr) {
    cta_channel_reduce_atomic(m, s1, red, sum);
    cta_channel_reduce_atomic(m, s2, red, sumsq);
  }
}

// dIn[n,iy,ix,c] = sum_{ky,kx} w[c,ky,kx] * dOut[n,(iy+pad-ky)/S,(ix+pad-kx)/S,c]; then (if the conv input was
// SiLU(BN(x_raw))) multiply by SiLU'(a*x_raw+b) and accumulate the BN-backward sums over dh and dh*xhat.
template <int K, int S>
__global__ void __launch_bounds__(CNN_THREADS, 2) dwconv_bwd_data_kernel(const bf16* __restrict__ dout, const bf16* __restrict__ w,
                                                                      const bf16* __restrict__ x_raw, const float* __restrict__ a_in,
                                                                      const float* __restrict__ b_in, const float* __restrict__ mean,
                                                                      const float* __restrict__ rstd, bf16* __restrict__ din,
                                                                      float* __restrict__ sums, int H, int W, int Ho, int Wo, int C) {
  extern __shared__ float dyn_sm[];
  float* wsm = dyn_sm;
  float* red = dyn_sm + K * K * DW_MAXV * 8;
  const ChanMap m = make_dw_map(C / 8, blockIdx.z);
  load_dw_weights<K>(wsm, w, m);
  constexpr int PAD = (K - 1) / 2;
  constexpr int NJ = (S == 1) ? (DW_PX + K - 1) : 4;
  const int n = blockIdx.y, XG = (W + DW_PX - 1) / DW_PX, items = H * XG;
  int ib, ie;
  cta_rows(items, ib, ie);
  float s1[8] = {0, 0, 0, 0, 0, 0, 0, 0}, s2[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (m.active) {
    const int cvl = threadIdx.x % m.c8;
    const bool has_bn = a_in != nullptr;
    float fa[8], fb[8], mu[8], rs[8];
    if (has_bn) {
      load8f(a_in + m.cv * 8, fa);
      load8f(b_in + m.cv * 8, fb);
      load8f(mean + m.cv * 8, mu);
      load8f(rstd + m.cv * 8, rs);
    }
    const bf16* g = dout + (size_t)n * Ho * Wo * C + (size_t)m.cv * 8;
    const size_t in_off = (size_t)n * H * W * C + (size_t)m.cv * 8;
    for (int it = ib + m.r0; it < ie; it += m.rows_per_pass) {
      const int iy = it / XG, ix0 = (it % XG) * DW_PX;
      float acc[DW_PX][8];
#pragma unroll
      for (int p = 0; p < DW_PX; ++p)
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[p][k] = 0.f;
#pragma unroll 1
      for (int ky = 0; ky < K; ++ky) {
        const int t = iy + PAD - ky;
        if (t < 0 || (t % S) != 0) continue;
        const int oy = t / S;
        if (oy >= Ho) continue;
        // candidate output columns: S == 1: ox = ix0 + PAD - (K-1) + j, kx = p + K-1 - j
        //                           S == 2: ox = ix0/2 - 1 + j,         kx = p + PAD + 2 - 2j   (ix0 is a multiple of 4)
        const int ox_base = (S == 1) ? (ix0 + PAD - (K - 1)) : (ix0 / 2 - 1);
        float wk[K][8];
        load_dw_row<K>(wsm, m.c8, cvl, ky, wk);
#pragma unroll
        for (int j = 0; j < NJ; ++j) {
          const int ox = ox_base + j;
          if (ox < 0 || ox >= Wo) continue;
          float v[8];
          load8bf(g + ((size_t)oy * Wo + ox) * C, v);
#pragma unroll
          for (int p = 0; p < DW_PX; ++p) {
            const int kx = (S == 1) ? (p + K - 1 - j) : (p + PAD + 2 - 2 * j);
            if (kx >= 0 && kx < K) {
#pragma unroll
              for (int k = 0; k < 8; ++k) acc[p][k] = fmaf(wk[kx][k], v[k], acc[p][k]);
            }
          }
        }
      }
#pragma unroll
      for (int p = 0; p < DW_PX; ++p) {
        const int ix = ix0 + p;
        if (ix >= W) continue;
        const size_t o = in_off + ((size_t)iy * W + ix) * C;
        if (has_bn) {
          float xr[8];
          load8bf(x_raw + o, xr);
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const float dh = acc[p][k] * silu_grad_f(fmaf(fa[k], xr[k], fb[k]));
            acc[p][k] = dh;
            s1[k] += dh;
            s2[k] += dh * (xr[k] - mu[k]) * rs[k];
          }
        }
        store8bf(din + o, acc[p]);
      }
    }
  }
  if (sums != nullptr) {
    cta_channel_reduce_atomic(m, s1, red, sums);
    cta_channel_reduce_atomic(m, s2, red, sums + C);
  }
}

// dw[ky][kx][c] += sum over (n, oy, ox) dOut * act(x); scratch layout [K*K][C] fp32.
// K x K taps x 8 channels do not fit in registers at a useful occupancy, so blockIdx.x % K splits the kernel
// rows over CTAs (ALLKY keeps the single-CTA variant for reference).
template <int K, int S>
__global__ void __launch_bounds__(CNN_THREADS, 2) dwconv_wgrad_kernel(const bf16* __restrict__ dout, const bf16* __restrict__ x,
                                                                      const float* __restrict__ a_in, const float* __restrict__ b_in,
                                                                      float* __restrict__ dw, int H, int W, int Ho, int Wo, int C) {
  __shared__ float red[CNN_THREADS * 8];
  constexpr bool ALLKY = false;   // all 9 taps of a 3x3 in one CTA was measured: 72 accumulators force 1 CTA/SM and it ran 17 % slower
  constexpr int NKY = ALLKY ? K : 1;
  // The K CTAs that share a tile (one per kernel row) are neighbours in blockIdx.x: they run at the same time and the
  // second..K-th read of dOut and x hits L2 (with ky on blockIdx.z they were a whole grid apart and every operand
  // streamed from HBM K times: ncu showed 796 MB of DRAM reads for 224 MB of data).
  const int ky0 = ALLKY ? 0 : blockIdx.x % K;
  const int xb = ALLKY ? blockIdx.x : blockIdx.x / K, xn = ALLKY ? gridDim.x : gridDim.x / K;
  const ChanMap m = make_dw_map(C / 8, blockIdx.z);
  constexpr int PAD = (K - 1) / 2, NIN = (DW_PX - 1) * S + K;
  const int n = blockIdx.y, XG = (Wo + DW_PX - 1) / DW_PX, items = Ho * XG;
  const int per = (items + xn - 1) / xn;
  const int ib = xb * per, ie = min(items, ib + per);
  float acc[NKY * K][8];
#pragma unroll
  for (int q = 0; q < NKY * K; ++q)
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[q][k] = 0.f;
  if (m.active) {
    const bool has_bn = a_in != nullptr;
    float fa[8], fb[8];
    if (has_bn) {
      load8f(a_in + m.cv * 8, fa);
      load8f(b_in + m.cv * 8, fb);
    }
    const bf16* xin = x + (size_t)n * H * W * C + (size_t)m.cv * 8;
    const bf16* g = dout + (size_t)n * Ho * Wo * C + (size_t)m.cv * 8;
    for (int it = ib + m.r0; it < ie; it += m.rows_per_pass) {
      const int oy = it / XG, ox0 = (it % XG) * DW_PX;
      float gd[DW_PX][8];
#pragma unroll
      for (int p = 0; p < DW_PX; ++p) {
        if (ox0 + p < Wo) load8bf(g + ((size_t)oy * Wo + ox0 + p) * C, gd[p]);
        else
#pragma unroll
          for (int k = 0; k < 8; ++k) gd[p][k] = 0.f;
      }
      const int ix_base = ox0 * S - PAD;
#pragma unroll
      for (int kyi = 0; kyi < NKY; ++kyi) {
        const int iy = oy * S - PAD + ky0 + kyi;
        if (iy < 0 || iy >= H) continue;
#pragma unroll
        for (int j = 0; j < NIN; ++j) {
          const int ix = ix_base + j;
          if (ix < 0 || ix >= W) continue;
          float v[8];
          load8bf(xin + ((size_t)iy * W + ix) * C, v);
          if (has_bn) {
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = silu_f(fmaf(fa[k], v[k], fb[k]));
          }
#pragma unroll
          for (int p = 0; p < DW_PX; ++p) {
            const int kx = j - p * S;
            if (kx >= 0 && kx < K) {
#pragma unroll
              for (int k = 0; k < 8; ++k) acc[kyi * K + kx][k] = fmaf(gd[p][k], v[k], acc[kyi * K + kx][k]);
            }
          }
        }
      }
    }
  }
#pragma unroll
  for (int kyi = 0; kyi < NKY; ++kyi)
#pragma unroll
    for (int kx = 0; kx < K; ++kx)
      cta_channel_reduce_atomic(m, acc[kyi * K + kx], red, dw + (size_t)((ky0 + kyi) * K + kx) * C);
}

// scratch [K*K][C] fp32 → weight gradient [C][K*K] bf16
__global__ void dw_v1_wgrad_cast_kernel(const float* __restrict__ scratch, bf16* __restrict__ dw, int C, int KK) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= C * KK) return;
  const int c = i / KK, t = i % KK;
  dw[i] = __float2bfloat16_rn(scratch[(size_t)t * C + c]);
}

static int dw_items_ctas(int N, int items) {
  int want = (148 * 8 + N - 1) / N;
  int maxc = (items + 15) / 16;
  if (want > maxc) want = maxc;
  return want < 1 ? 1 : want;
}

#define DW_DISPATCH(KERNEL, ...)                                                                     \
  do {                                                                                               \
    if (k == 3 && stride == 1) KERNEL<3, 1> __VA_ARGS__;                                              \
    else if (k == 3 && stride == 2) KERNEL<3, 2> __VA_ARGS__;                                         \
    else if (k == 5 && stride == 1) KERNEL<5, 1> __VA_ARGS__;                                         \
    else if (k == 5 && stride == 2) KERNEL<5, 2> __VA_ARGS__;                                         \
    else RT1_REQUIRE(false, "depthwise conv: unsupported kernel %d stride %d", k, stride);            \
  } while (0)

// x [N,H,W,C] raw (a_in/b_in = producer BN scale/shift, SiLU applied on load) or already-activated (a_in = null);
// y [N,Ho,Wo,C] raw output; sum/sumsq fp32 [C] (caller-zeroed) or null.
RT1_API int rt1_dwconv_v1_fwd(const void* x, const float* a_in, const float* b_in, const void* w, void* y, float* sum, float* sumsq,
                           int N, int H, int W, int C, int k, int stride, cudaStream_t stream) {
  RT1_REQUIRE(C % 8 == 0, "rt1_dwconv_v1_fwd: C=%d must be a multiple of 8", C);
  const int pad = (k - 1) / 2, Ho = (H + 2 * pad - k) / stride + 1, Wo = (W + 2 * pad - k) / stride + 1;
  const int items = Ho * ((Wo + DW_PX - 1) / DW_PX);
  dim3 grid(dw_items_ctas(N, items), N, dw_chunks(C));
  const size_t smem = ((size_t)k * k * DW_MAXV * 8 + CNN_THREADS * 8) * sizeof(float);
  static bool attr_set = false;
  if (!attr_set) {
    const int mx = (25 * DW_MAXV * 8 + CNN_THREADS * 8) * 4;
    RT1_CHECK_CUDA(cudaFuncSetAttribute(dwconv_fwd_kernel<5, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
    RT1_CHECK_CUDA(cudaFuncSetAttribute(dwconv_fwd_kernel<5, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
    RT1_CHECK_CUDA(cudaFuncSetAttribute(dwconv_fwd_kernel<3, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
    RT1_CHECK_CUDA(cudaFuncSetAttribute(dwconv_fwd_kernel<3, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
    attr_set = true;
  }
  DW_DISPATCH(dwconv_fwd_kernel, <<<grid, CNN_THREADS, smem, stream>>>((const bf16*)x, a_in, b_in, (const bf16*)w, (bf16*)y, sum, sumsq, H, W, Ho, Wo, C));
  return rt1_check_launch("rt1_dwconv_v1_fwd");
}

// dout [N,Ho,Wo,C]; x_raw [N,H,W,C] + a_in/b_in/mean/rstd (or all null when the conv input was not BN+SiLU);
// din [N,H,W,C]: d/d(x_raw-normalised pre-activation) "dh" when BN is present, else d/d(input);
// sums fp32 [2,C] (caller-zeroed) receives sum dh and sum dh*xhat (BN case).
RT1_API int rt1_dwconv_v1_bwd_data(const void* dout, const void* w, const void* x_raw, const float* a_in, const float* b_in,
                                const float* mean, const float* rstd, void* din, float* sums, int N, int H, int W, int C, int k,
                                int stride, cudaStream_t stream) {
  RT1_REQUIRE(C % 8 == 0, "rt1_dwconv_v1_bwd_data: C=%d must be a multiple of 8", C);
  const int pad = (k - 1) / 2, Ho = (H + 2 * pad - k) / stride + 1, Wo = (W + 2 * pad - k) / stride + 1;
  const int items = H * ((W + DW_PX - 1) / DW_PX);
  dim3 grid(dw_items_ctas(N, items), N, dw_chunks(C));
  const size_t smem = ((size_t)k * k * DW_MAXV * 8 + CNN_THREADS * 8) * sizeof(float);
  static bool attr_set = false;
  if (!attr_set) {
    const int mx = (25 * DW_MAXV * 8 + CNN_THREADS * 8) * 4;
    RT1_CHECK_CUDA(cudaFuncSetAttribute(dwconv_bwd_data_kernel<5, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
    RT1_CHECK_CUDA(cudaFuncSetAttribute(dwconv_bwd_data_kernel<5, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
    RT1_CHECK_CUDA(cudaFuncSetAttribute(dwconv_bwd_data_kernel<3, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
    RT1_CHECK_CUDA(cudaFuncSetAttribute(dwconv_bwd_data_kernel<3, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
    attr_set = true;
  }
  DW_DISPATCH(dwconv_bwd_data_kernel, <<<grid, CNN_THREADS, smem, stream>>>((const bf16*)dout, (const bf16*)w, (const bf16*)x_raw, a_in, b_in, mean, rstd, (bf16*)din, sums, H, W, Ho, Wo, C));
  return rt1_check_launch("rt1_dwconv_v1_bwd_data");
}

// scratch: fp32 [k*k*C], zeroed by this call; dw: bf16 [C, k, k]
RT1_API int rt1_dwconv_v1_wgrad(const void* dout, const void* x, const float* a_in, const float* b_in, float* scratch, void* dw, int N,
                             int H, int W, int C, int k, int stride, cudaStream_t stream) {
  RT1_REQUIRE(C % 8 == 0, "rt1_dwconv_v1_wgrad: C=%d must be a multiple of 8", C);
  const int pad = (k - 1) / 2, Ho = (H + 2 * pad - k) / stride + 1, Wo = (W + 2 * pad - k) / stride + 1;
  const int items = Ho * ((Wo + DW_PX - 1) / DW_PX);
  RT1_CHECK_CUDA(cudaMemsetAsync(scratch, 0, sizeof(float) * (size_t)k * k * C, stream));
  int ctas = dw_items_ctas(N, items);
  const int zsplit = k;   // kernel rows split over blockIdx.z
  ctas = (ctas + zsplit - 1) / zsplit;
  dim3 grid((ctas < 1 ? 1 : ctas) * zsplit, N, dw_chunks(C));
  DW_DISPATCH(dwconv_wgrad_kernel, <<<grid, CNN_THREADS, 0, stream>>>((const bf16*)dout, (const bf16*)x, a_in, b_in, scratch, H, W, Ho, Wo, C));
  int rc = rt1_check_launch("rt1_dwconv_v1_wgrad");
  if (rc) return rc;
  dw_v1_wgrad_cast_kernel<<<rt1_cdiv((long)C * k * k, 256), 256, 0, stream>>>(scratch, (bf16*)dw, C, k * k);
  return rt1_check_launch("rt1_dwconv_v1_wgrad(cast)");
}

