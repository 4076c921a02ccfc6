// LayerNorm forward/backward over bf16 rows (decoder pre-LN; replaces ATen layer_norm in K11).
//
// Forward: one warp per row, the row lives in registers (D <= 2048), two-pass mean/variance in fp32, writes
// y (bf16) and per-row mean / rstd (fp32) for the backward.
// Backward: one warp per row for dx; per-column dgamma/dbeta partials are accumulated per CTA in shared
// memory over the rows that CTA owns and written to a [grid, 2, D] fp32 scratch, which a second kernel folds
// and stores as bf16 straight into the parameter-gradient slots.
#include "common.cuh"

#define LN_MAX_PER_LANE 64  // D <= 2048

template <int VPL>  // 8-element vectors per lane
__global__ void __launch_bounds__(256) ln_fwd_kernel(const bf16* __restrict__ x, const bf16* __restrict__ w,
                                                     const bf16* __restrict__ b, bf16* __restrict__ y,
                                                     float* __restrict__ mean_out, float* __restrict__ rstd_out, int rows,
                                                     int D, float eps) {
  pdl_enter();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
  const int row = blockIdx.x * wpb + warp;
  if (row >= rows) return;
  const bf16x8* xr = reinterpret_cast<const bf16x8*>(x + (size_t)row * D);
  float v[VPL * 8];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    unpack8(xr[lane + 32 * i], v + 8 * i);
#pragma unroll
    for (int k = 0; k < 8; ++k) s += v[8 * i + k];
  }
  const float mean = warp_sum(s) / D;
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < VPL * 8; ++i) {
    float d = v[i] - mean;
    q += d * d;
  }
  const float rstd = rsqrtf(warp_sum(q) / D + eps);
  bf16x8* yr = reinterpret_cast<bf16x8*>(y + (size_t)row * D);
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    float wv[8], bv[8], o[8];
    unpack8(reinterpret_cast<const bf16x8*>(w)[lane + 32 * i], wv);
    unpack8(reinterpret_cast<const bf16x8*>(b)[lane + 32 * i], bv);
#pragma unroll
    for (int k = 0; k < 8; ++k) o[k] = (v[8 * i + k] - mean) * rstd * wv[k] + bv[k];
    yr[lane + 32 * i] = pack8(o);
  }
  if (lane == 0) {
    if (mean_out) mean_out[row] = mean;
    if (rstd_out) rstd_out[row] = rstd;
  }
}

template <int VPL>
__global__ void __launch_bounds__(256) ln_bwd_kernel(const bf16* __restrict__ x, const bf16* __restrict__ w,
                                                     const bf16* __restrict__ dy, const float* __restrict__ mean_in,
                                                     const float* __restrict__ rstd_in, bf16* __restrict__ dx,
                                                     float* __restrict__ partial, int rows, int D) {
  pdl_enter();
  extern __shared__ float sm[];  // [2][D] per-CTA dgamma / dbeta partials
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
  for (int i = threadIdx.x; i < 2 * D; i += blockDim.x) sm[i] = 0.f;
  __syncthreads();
  float wv[VPL * 8];
#pragma unroll
  for (int i = 0; i < VPL; ++i) unpack8(reinterpret_cast<const bf16x8*>(w)[lane + 32 * i], wv + 8 * i);
  float dg[VPL * 8], db[VPL * 8];
#pragma unroll
  for (int i = 0; i < VPL * 8; ++i) dg[i] = db[i] = 0.f;
  for (int row = blockIdx.x * wpb + warp; row < rows; row += gridDim.x * wpb) {
    const bf16x8* xr = reinterpret_cast<const bf16x8*>(x + (size_t)row * D);
    const bf16x8* gr = reinterpret_cast<const bf16x8*>(dy + (size_t)row * D);
    const float mean = mean_in[row], rstd = rstd_in[row];
    float xh[VPL * 8], g[VPL * 8];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
      unpack8(xr[lane + 32 * i], xh + 8 * i);
      unpack8(gr[lane + 32 * i], g + 8 * i);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int j = 8 * i + k;
        xh[j] = (xh[j] - mean) * rstd;
        dg[j] += g[j] * xh[j];
        db[j] += g[j];
        g[j] *= wv[j];
        s1 += g[j];
        s2 += g[j] * xh[j];
      }
    }
    s1 = warp_sum(s1) / D;
    s2 = warp_sum(s2) / D;
    bf16x8* dr = reinterpret_cast<bf16x8*>(dx + (size_t)row * D);
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
      float o[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) o[k] = rstd * (g[8 * i + k] - s1 - xh[8 * i + k] * s2);
      dr[lane + 32 * i] = pack8(o);
    }
  }
#pragma unroll
  for (int i = 0; i < VPL; ++i)
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int col = (lane + 32 * i) * 8 + k;
      atomicAdd(&sm[col], dg[8 * i + k]);
      atomicAdd(&sm[D + col], db[8 * i + k]);
    }
  __syncthreads();
  for (int i = threadIdx.x; i < 2 * D; i += blockDim.x) partial[(size_t)blockIdx.x * 2 * D + i] = sm[i];
}

// dgamma / dbeta = sums of the per-CTA partials [nblk][2][D].  32 columns x 8 partial lanes per CTA, four loads in flight per
// thread (one thread per column walking all ~148 partials with dependent loads took 12.7 us for D = 512).
__global__ void __launch_bounds__(256) ln_param_grad_kernel(const float* __restrict__ partial, int nblk, int D, bf16* __restrict__ dw,
                                                            bf16* __restrict__ db) {
  pdl_enter();
  __shared__ float sa[8][33], sc[8][33];
  const int cl = threadIdx.x & 31, col = blockIdx.x * 32 + cl, l0 = threadIdx.x >> 5;
  float a0 = 0.f, a1 = 0.f, c0 = 0.f, c1 = 0.f;
  if (col < D) {
    int i = l0;
    for (; i + 8 < nblk; i += 16) {
      const float* p0 = partial + (size_t)i * 2 * D + col;
      const float* p1 = partial + (size_t)(i + 8) * 2 * D + col;
      const float x0 = p0[0], y0 = p0[D], x1 = p1[0], y1 = p1[D];
      a0 += x0; c0 += y0; a1 += x1; c1 += y1;
    }
    for (; i < nblk; i += 8) {
      a0 += partial[(size_t)i * 2 * D + col];
      c0 += partial[(size_t)i * 2 * D + D + col];
    }
  }
  sa[l0][cl] = a0 + a1;
  sc[l0][cl] = c0 + c1;
  __syncthreads();
  if (l0 == 0 && col < D) {
    float a = 0.f, c = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      a += sa[k][cl];
      c += sc[k][cl];
    }
    dw[col] = __float2bfloat16_rn(a);
    db[col] = __float2bfloat16_rn(c);
  }
}

RT1_API int rt1_layernorm_fwd(const void* x, const void* w, const void* b, void* y, float* mean, float* rstd, int rows,
                              int D, float eps, cudaStream_t stream) {
  RT1_REQUIRE(D % 256 == 0 && D <= 2048, "rt1_layernorm_fwd: D=%d must be a multiple of 256 and <= 2048", D);
  const int wpb = 8, grid = rt1_cdiv(rows, wpb);
#define LN_F(V) rt1_launch(ln_fwd_kernel<V>, grid, wpb * 32, 0, stream, (const bf16*)x, (const bf16*)w, (const bf16*)b, (bf16*)y, mean, rstd, rows, D, eps)
  switch (D / 256) {
    case 1: LN_F(1); break;
    case 2: LN_F(2); break;
    case 4: LN_F(4); break;
    case 8: LN_F(8); break;
    default: RT1_REQUIRE(false, "rt1_layernorm_fwd: unsupported D=%d", D);
  }
#undef LN_F
  return rt1_check_launch("rt1_layernorm_fwd");
}

// scratch: fp32 [rt1_layernorm_bwd_scratch_floats(rows, D)]
RT1_API long long rt1_layernorm_bwd_scratch_floats(int rows, int D) {
  int grid = rt1_cdiv(rows, 8);
  if (grid > 148) grid = 148;
  return (long long)grid * 2 * D;
}

RT1_API int rt1_layernorm_bwd(const void* x, const void* w, const void* dy, const float* mean, const float* rstd,
                              void* dx, void* dw, void* db, float* scratch, int rows, int D, cudaStream_t stream) {
  RT1_REQUIRE(D % 256 == 0 && D <= 2048, "rt1_layernorm_bwd: D=%d must be a multiple of 256 and <= 2048", D);
  int grid = rt1_cdiv(rows, 8);
  if (grid > 148) grid = 148;
  const size_t smem = 2 * (size_t)D * sizeof(float);
#define LN_B(V) rt1_launch(ln_bwd_kernel<V>, grid, 256, smem, stream, (const bf16*)x, (const bf16*)w, (const bf16*)dy, mean, rstd, (bf16*)dx, scratch, rows, D)
  switch (D / 256) {
    case 1: LN_B(1); break;
    case 2: LN_B(2); break;
    case 4: LN_B(4); break;
    case 8: LN_B(8); break;
    default: RT1_REQUIRE(false, "rt1_layernorm_bwd: unsupported D=%d", D);
  }
#undef LN_B
  int rc = rt1_check_launch("rt1_layernorm_bwd");
  if (rc) return rc;
  rt1_launch(ln_param_grad_kernel, rt1_cdiv(D, 32), 256, 0, stream, scratch, grid, D, (bf16*)dw, (bf16*)db);
  return rt1_check_launch("rt1_layernorm_param_grad");
}

// ----------------------------------------------------------------------------------------------------------------
// rt1_layernorm_bwd_noalloc — LayerNorm backward for hosts that cannot hand us a scratch buffer.
//
// Who this is for.  rt1_layernorm_bwd() needs a caller-provided fp32 scratch for the per-CTA dgamma/dbeta partial
// sums.  Some embedders of this library have no way to supply one: a robot-control process that reserved all of its
// device memory at start-up and runs fine-tuning steps inside a loop where cudaMalloc is off limits (it synchronises
// the whole device and may fail late, mid-episode), or a host that has given every free byte to another allocator
// (a framework's caching pool, an MPS partition with a fixed budget) so that a raw cudaMalloc from a plug-in simply
// returns out-of-memory.  For them the scratch lives inside the library image itself, below, as a fixed-size
// __device__ array: it is part of the module every GPU loads, costs LN_NOALLOC_MAX_FLOATS * 4 B = 1.2 MB per GPU
// until the process exits, and needs no allocation call of any kind.  This entry point calls no cudaMalloc / cudaFree
// / memset-allocating or other allocating routine; inputs and outputs are still the caller's.
//
// Limits.  D (row width) must be a multiple of 256 and at most LN_NOALLOC_MAX_D = 1024; any number of rows (the
// reduction grid is capped at 148 CTAs, so the scratch need is 148 * 2 * D floats at most).  Larger D is rejected with
// an error message, not truncated.
//
// Set-up.  In host code the name of a __device__ array is only a host-side handle, not a device address — passing
// it to a kernel would fault or touch the wrong memory — so the device address is looked up with
// cudaGetSymbolAddress.  That happens on the FIRST call made on each device (with that device current), and the
// result is cached per device.  Make that first call on every device you will use before starting any stream capture.
//
// Sharing.  Every call on a device uses the same array and finds the previous call's contents in it.  Nothing in it
// has to start at zero: each CTA overwrites its own [2 * D] slice completely and the fold kernel reads exactly the
// slices written by this launch, so no clearing is done.  Two calls whose kernels overlap on the GPU (different
// streams, or the same call issued from two threads) race on the array and give wrong dgamma/dbeta — serialise them
// (same stream, or an event between them).  rt1_layernorm_bwd() with caller scratch has no such restriction.
#include <mutex>

#define LN_NOALLOC_MAX_D 1024
#define LN_NOALLOC_MAX_GRID 148
#define LN_NOALLOC_MAX_FLOATS (LN_NOALLOC_MAX_GRID * 2 * LN_NOALLOC_MAX_D)
#define LN_NOALLOC_MAX_DEVICES 64

static __device__ __align__(256) float g_ln_noalloc_scratch[LN_NOALLOC_MAX_FLOATS];

static float* g_ln_noalloc_ptr[LN_NOALLOC_MAX_DEVICES] = {nullptr};
static std::mutex g_ln_noalloc_mutex;

// Device address of the scratch array on the CURRENT device; resolved once per device.
static int ln_noalloc_scratch_ptr(float** out) {
  int dev = -1;
  RT1_CHECK_CUDA(cudaGetDevice(&dev));
  RT1_REQUIRE(dev >= 0 && dev < LN_NOALLOC_MAX_DEVICES, "rt1_layernorm_bwd_noalloc: device index %d not supported", dev);
  std::lock_guard<std::mutex> lock(g_ln_noalloc_mutex);
  if (g_ln_noalloc_ptr[dev] == nullptr) {
    float* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&p, g_ln_noalloc_scratch);
    if (e != cudaSuccess || p == nullptr) {
      rt1_set_error("rt1_layernorm_bwd_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
      return e != cudaSuccess ? (int)e : -1;
    }
    g_ln_noalloc_ptr[dev] = p;
  }
  *out = g_ln_noalloc_ptr[dev];
  return 0;
}

RT1_API long long rt1_layernorm_bwd_noalloc_max_d(void) { return LN_NOALLOC_MAX_D; }

RT1_API int rt1_layernorm_bwd_noalloc(const void* x, const void* w, const void* dy, const float* mean, const float* rstd,
                                      void* dx, void* dw, void* db, int rows, int D, cudaStream_t stream) {
  RT1_REQUIRE(rows > 0, "rt1_layernorm_bwd_noalloc: rows=%d must be positive", rows);
  RT1_REQUIRE(D % 256 == 0 && D > 0, "rt1_layernorm_bwd_noalloc: D=%d must be a positive multiple of 256", D);
  RT1_REQUIRE(D <= LN_NOALLOC_MAX_D,
              "rt1_layernorm_bwd_noalloc: D=%d exceeds the built-in scratch (max D %d); use rt1_layernorm_bwd with caller scratch",
              D, LN_NOALLOC_MAX_D);
  int grid = rt1_cdiv(rows, 8);
  if (grid > LN_NOALLOC_MAX_GRID) grid = LN_NOALLOC_MAX_GRID;
  RT1_REQUIRE((long long)grid * 2 * D <= LN_NOALLOC_MAX_FLOATS, "rt1_layernorm_bwd_noalloc: scratch need exceeds the built-in array");
  float* scratch = nullptr;
  int rc = ln_noalloc_scratch_ptr(&scratch);
  if (rc) return rc;
  const size_t smem = 2 * (size_t)D * sizeof(float);
  switch (D / 256) {
    case 1: rt1_launch(ln_bwd_kernel<1>, grid, 256, smem, stream, (const bf16*)x, (const bf16*)w, (const bf16*)dy, mean, rstd, (bf16*)dx, scratch, rows, D); break;
    case 2: rt1_launch(ln_bwd_kernel<2>, grid, 256, smem, stream, (const bf16*)x, (const bf16*)w, (const bf16*)dy, mean, rstd, (bf16*)dx, scratch, rows, D); break;
    case 4: rt1_launch(ln_bwd_kernel<4>, grid, 256, smem, stream, (const bf16*)x, (const bf16*)w, (const bf16*)dy, mean, rstd, (bf16*)dx, scratch, rows, D); break;
    default: RT1_REQUIRE(false, "rt1_layernorm_bwd_noalloc: unsupported D=%d (supported: 256, 512, 1024)", D);
  }
  rc = rt1_check_launch("rt1_layernorm_bwd_noalloc");
  if (rc) return rc;
  rt1_launch(ln_param_grad_kernel, rt1_cdiv(D, 32), 256, 0, stream, scratch, grid, D, (bf16*)dw, (bf16*)db);
  return rt1_check_launch("rt1_layernorm_bwd_noalloc(param_grad)");
}
