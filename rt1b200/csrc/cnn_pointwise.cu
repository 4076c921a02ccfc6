// Elementwise / reduction kernels of the FiLM-EfficientNet pipeline in NHWC bf16 (K3–K8).
//
// BatchNorm in training mode needs batch statistics before it can be applied, so an MBConv block cannot be one
// kernel; instead every producer (GEMM / depthwise kernel) emits per-channel sum / sum-of-squares from its
// epilogue, and the BN apply (+SiLU, +SE scale, +FiLM, +residual) is folded into whichever pass reads the raw
// tensor next.  The kernels here are those passes:
//   bn_finalize            stats → scale/shift (+ running statistics), one tiny launch per BN layer
//   bn_act_pool            SE squeeze: per-(frame, channel) sum of SiLU(BN(x))
//   bn_act_scale           y = SiLU(BN(x)) * s[n, c]           (SE excite; s = null → plain BN+SiLU)
//   block_finalize         out = (1 + gamma) * (skip + sd * BN(y)) + beta       (BN3 + stochastic depth + residual + FiLM)
//   block_finalize_bwd_*   the matching backward (per-frame sums → FiLM / BN3 gradients → dY, dskip)
//   bn_act_bwd_*           backward of y = act(BN(x)) [* s + pooled-gradient], reduce + apply passes
// Every kernel uses the channel-stationary mapping of cnn_common.cuh; grid = (ctas_per_frame, N, channel_chunks).
#include "cnn_common.cuh"

// ------------------------------------------------------------------------------------------------ bn_finalize
__global__ void bn_finalize_kernel(const float* __restrict__ sum, const float* __restrict__ sumsq, float count,
                                   const bf16* __restrict__ gamma, const bf16* __restrict__ beta, float* __restrict__ running_mean,
                                   float* __restrict__ running_var, float momentum, float eps, int training, float* __restrict__ a,
                                   float* __restrict__ b, float* __restrict__ mean_out, float* __restrict__ rstd_out, int C) {
  pdl_enter();
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  float mean, var;
  if (training) {
    mean = sum[c] / count;
    var = fmaxf(sumsq[c] / count - mean * mean, 0.f);
    if (running_mean != nullptr) {
      running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
      const float unbiased = count > 1.f ? var * count / (count - 1.f) : var;
      running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
    }
  } else {
    mean = running_mean[c];
    var = running_var[c];
  }
  const float rstd = rsqrtf(var + eps);
  const float g = __bfloat162float(gamma[c]);
  a[c] = g * rstd;
  b[c] = __bfloat162float(beta[c]) - mean * g * rstd;
  mean_out[c] = mean;
  rstd_out[c] = rstd;
}

RT1_API int rt1_bn_finalize(const float* sum, const float* sumsq, float count, const void* gamma, const void* beta,
                            float* running_mean, float* running_var, float momentum, float eps, int training, float* a, float* b,
                            float* mean, float* rstd, int C, cudaStream_t stream) {
  rt1_launch(bn_finalize_kernel, rt1_cdiv(C, 128), 128, 0, stream, sum, sumsq, count, (const bf16*)gamma, (const bf16*)beta, running_mean,
                                                          running_var, momentum, eps, training, a, b, mean, rstd, C);
  return rt1_check_launch("rt1_bn_finalize");
}

// --------------------------------------------------------------------------------------------- bn_act_pool (SE squeeze)
__global__ void __launch_bounds__(CNN_THREADS) bn_act_pool_kernel(const bf16* __restrict__ x, const float* __restrict__ a,
                                                                  const float* __restrict__ b, float* __restrict__ pooled, int HW,
                                                                  int C) {
  pdl_enter();
  __shared__ float sm[CNN_THREADS * 8];
  const ChanMap m = make_chan_map(C / 8);
  const int n = blockIdx.y;
  int rb, re;
  cta_rows(HW, rb, re);
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (m.active) {
    float fa[8], fb[8];
    load8f(a + m.cv * 8, fa);
    load8f(b + m.cv * 8, fb);
    const bf16* base = x + ((size_t)n * HW) * C + (size_t)m.cv * 8;
    for (int r = rb + m.r0; r < re; r += m.rows_per_pass) {
      float f[8];
      load8bf(base + (size_t)r * C, f);
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] += silu_f(fmaf(fa[k], f[k], fb[k]));
    }
  }
  cta_channel_reduce_atomic(m, acc, sm, pooled + (size_t)n * C);
}

// pooled [N, C] fp32 must be zeroed by the caller; receives SUMS over the H*W positions
RT1_API int rt1_bn_act_pool(const void* x, const float* a, const float* b, float* pooled, int N, int HW, int C,
                            cudaStream_t stream) {
  RT1_REQUIRE(C % 8 == 0, "rt1_bn_act_pool: C=%d must be a multiple of 8", C);
  dim3 grid(cnn_ctas_per_frame(N, HW, C), N, cnn_chan_chunks(C));
  rt1_launch(bn_act_pool_kernel, grid, CNN_THREADS, 0, stream, (const bf16*)x, a, b, pooled, HW, C);
  return rt1_check_launch("rt1_bn_act_pool");
}

// ------------------------------------------------------------------------------------------------ bn_act_scale
// y = act(a*x + b) * s[n, c]; act = SiLU when `silu`, identity otherwise; s may be null.
__global__ void __launch_bounds__(CNN_THREADS) bn_act_scale_kernel(const bf16* __restrict__ x, const float* __restrict__ a,
                                                                   const float* __restrict__ b, const float* __restrict__ s,
                                                                   bf16* __restrict__ y, int HW, int C, int silu) {
  pdl_enter();
  const ChanMap m = make_chan_map(C / 8);
  if (!m.active) return;
  const int n = blockIdx.y;
  int rb, re;
  cta_rows(HW, rb, re);
  p2 fa[4], fb[4], fs[4];
  load8p(a + m.cv * 8, fa);
  load8p(b + m.cv * 8, fb);
  if (s != nullptr) load8p(s + (size_t)n * C + m.cv * 8, fs);
  else
#pragma unroll
    for (int j = 0; j < 4; ++j) fs[j] = p2_pack(1.f, 1.f);
  const size_t off = ((size_t)n * HW) * C + (size_t)m.cv * 8;
  // two rows in flight per thread; channel pairs in packed fp32
  for (int r = rb + m.r0; r < re; r += 2 * m.rows_per_pass) {
    const int r2 = r + m.rows_per_pass;
    const bool has2 = r2 < re;
    bf16x8 raw[2];
    raw[0] = *reinterpret_cast<const bf16x8*>(x + off + (size_t)r * C);
    if (has2) raw[1] = *reinterpret_cast<const bf16x8*>(x + off + (size_t)r2 * C);
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      if (h == 1 && !has2) break;
      p2 f[4];
      float o[8];
      unpack8p(raw[h], f);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const p2 u = p2_fma(fa[j], f[j], fb[j]);
        const p2 yv = silu ? p2_mul(p2_mul(u, p2_sigmoid(u)), fs[j]) : p2_mul(u, fs[j]);
        p2_unpack(yv, o[2 * j], o[2 * j + 1]);
      }
      store8bf(y + off + (size_t)(h ? r2 : r) * C, o);
    }
  }
}

RT1_API int rt1_bn_act_scale(const void* x, const float* a, const float* b, const float* s, void* y, int N, int HW, int C, int silu,
                             cudaStream_t stream) {
  RT1_REQUIRE(C % 8 == 0, "rt1_bn_act_scale: C=%d must be a multiple of 8", C);
  dim3 grid(cnn_ctas_per_frame(N, HW, C), N, cnn_chan_chunks(C));
  rt1_launch(bn_act_scale_kernel, grid, CNN_THREADS, 0, stream, (const bf16*)x, a, b, s, (bf16*)y, HW, C, silu);
  return rt1_check_launch("rt1_bn_act_scale");
}

// ---------------------------------------------------------------------------------------------- block_finalize
// out = (1 + gamma[n,c]) * (skip + sd[n] * (a*y + b)) + beta[n,c]
//   a/b null  → no BatchNorm (u = y);   skip null → no residual;   sd null → 1;   gb null → no FiLM
//   gb: bf16 [N, gb_stride]; beta at gb[n, gb_off + c], gamma at gb[n, gb_off + C + c]
__global__ void __launch_bounds__(CNN_THREADS) block_finalize_kernel(const bf16* __restrict__ y, const float* __restrict__ a,
                                                                     const float* __restrict__ b, const bf16* __restrict__ skip,
                                                                     const float* __restrict__ sd, const bf16* __restrict__ gb,
                                                                     long gb_stride, long gb_off, bf16* __restrict__ out, int HW, int C) {
  pdl_enter();
  const ChanMap m = make_chan_map(C / 8);
  if (!m.active) return;
  const int n = blockIdx.y;
  int rb, re;
  cta_rows(HW, rb, re);
  float fa[8], fb[8], g1[8], be[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    fa[k] = 1.f; fb[k] = 0.f; g1[k] = 1.f; be[k] = 0.f;
  }
  const float sdn = sd != nullptr ? sd[n] : 1.f;
  if (a != nullptr) {
    load8f(a + m.cv * 8, fa);
    load8f(b + m.cv * 8, fb);
  }
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    fa[k] *= sdn;
    fb[k] *= sdn;
  }
  if (gb != nullptr) {
    float t[8];
    load8bf(gb + (size_t)n * gb_stride + gb_off + m.cv * 8, be);
    load8bf(gb + (size_t)n * gb_stride + gb_off + C + m.cv * 8, t);
#pragma unroll
    for (int k = 0; k < 8; ++k) g1[k] = 1.f + t[k];
  }
  const size_t off = ((size_t)n * HW) * C + (size_t)m.cv * 8;
  for (int r = rb + m.r0; r < re; r += m.rows_per_pass) {
    float f[8], sk[8];
    load8bf(y + off + (size_t)r * C, f);
    if (skip != nullptr) load8bf(skip + off + (size_t)r * C, sk);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float z = fmaf(fa[k], f[k], fb[k]);
      if (skip != nullptr) z += sk[k];
      f[k] = fmaf(g1[k], z, be[k]);
    }
    store8bf(out + off + (size_t)r * C, f);
  }
}

RT1_API int rt1_block_finalize(const void* y, const float* a, const float* b, const void* skip, const float* sd, const void* gb,
                               long long gb_stride, long long gb_off, void* out, int N, int HW, int C, cudaStream_t stream) {
  RT1_REQUIRE(C % 8 == 0, "rt1_block_finalize: C=%d must be a multiple of 8", C);
  dim3 grid(cnn_ctas_per_frame(N, HW, C), N, cnn_chan_chunks(C));
  rt1_launch(block_finalize_kernel, grid, CNN_THREADS, 0, stream, (const bf16*)y, a, b, (const bf16*)skip, sd, (const bf16*)gb, gb_stride,
                                                          gb_off, (bf16*)out, HW, C);
  return rt1_check_launch("rt1_block_finalize");
}

// ------------------------------------------------------------------------------------- block_finalize backward
// Pass 1: per-(frame, channel) sums  S1 = sum dOut, S2 = sum dOut * skip, S3 = sum dOut * yhat
//   yhat = (y - mean) * rstd when BN is present, else yhat = y.
__global__ void __launch_bounds__(CNN_THREADS) block_finalize_bwd_reduce_kernel(const bf16* __restrict__ dout, const bf16* __restrict__ y,
                                                                                const bf16* __restrict__ skip, const float* __restrict__ mean,
                                                                                const float* __restrict__ rstd, float* __restrict__ S, int HW,
                                                                                int C, int N) {
  pdl_enter();
  __shared__ float sm[CNN_THREADS * 8];
  const ChanMap m = make_chan_map(C / 8);
  const int n = blockIdx.y;
  int rb, re;
  cta_rows(HW, rb, re);
  float s1[8] = {0, 0, 0, 0, 0, 0, 0, 0}, s2[8] = {0, 0, 0, 0, 0, 0, 0, 0}, s3[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (m.active) {
    float mu[8], rs[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      mu[k] = 0.f;
      rs[k] = 1.f;
    }
    if (mean != nullptr) {
      load8f(mean + m.cv * 8, mu);
      load8f(rstd + m.cv * 8, rs);
    }
    const size_t off = ((size_t)n * HW) * C + (size_t)m.cv * 8;
    for (int r = rb + m.r0; r < re; r += m.rows_per_pass) {
      float g[8], f[8], sk[8];
      load8bf(dout + off + (size_t)r * C, g);
      load8bf(y + off + (size_t)r * C, f);
      if (skip != nullptr) load8bf(skip + off + (size_t)r * C, sk);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        s1[k] += g[k];
        if (skip != nullptr) s2[k] += g[k] * sk[k];
        s3[k] += g[k] * (f[k] - mu[k]) * rs[k];
      }
    }
  }
  const size_t plane = (size_t)N * C;
  cta_channel_reduce_atomic(m, s1, sm, S + (size_t)n * C);
  cta_channel_reduce_atomic(m, s2, sm, S + plane + (size_t)n * C);
  cta_channel_reduce_atomic(m, s3, sm, S + 2 * plane + (size_t)n * C);
}

// Pass 2 (tiny): FiLM gradients per (n, c), BN3 parameter gradients per c, and the two BN-backward constants.
//   dbeta_film = S1;  dgamma_film = S2 + sd * (g3 * S3 + b3 * S1)      [with BN: u = g3*yhat + b3; without: u = y → S3]
//   dbeta3 = sum_n sd (1+gamma) S1;  dgamma3 = sum_n sd (1+gamma) S3;  c1 = dbeta3 / M;  c2 = dgamma3 / M
// block (32 channels, 8 frame lanes): a thread walks frames ny, ny + 8, ... so the per-frame loads of one channel are spread
// over 8 threads instead of forming one 48-deep dependent chain (31 us per launch before, for a few KB of data)
__global__ void __launch_bounds__(256) block_finalize_bwd_combine_kernel(const float* __restrict__ S, const bf16* __restrict__ gb, long gb_stride, long gb_off,
                                                  const float* __restrict__ sd, const bf16* __restrict__ gamma3,
                                                  const bf16* __restrict__ beta3, bf16* __restrict__ dgb, bf16* __restrict__ dgamma3,
                                                  bf16* __restrict__ dbeta3, float* __restrict__ c1, float* __restrict__ c2, int N, int C,
                                                  float inv_count) {
  pdl_enter();
  __shared__ float sdb[8][33], sdg[8][33];
  const int tx = threadIdx.x, ny = threadIdx.y;
  const int c = blockIdx.x * 32 + tx;
  const size_t plane = (size_t)N * C;
  const bool has_bn = gamma3 != nullptr;
  float db = 0.f, dg = 0.f;
  if (c < C) {
    const float g3 = has_bn ? __bfloat162float(gamma3[c]) : 1.f, b3 = has_bn ? __bfloat162float(beta3[c]) : 0.f;
#pragma unroll 4
    for (int n = ny; n < N; n += 8) {
      const float s1 = S[(size_t)n * C + c], s2 = S[plane + (size_t)n * C + c], s3 = S[2 * plane + (size_t)n * C + c];
      const float sdn = sd != nullptr ? sd[n] : 1.f;
      float gam = 0.f;
      if (gb != nullptr) {
        gam = __bfloat162float(gb[(size_t)n * gb_stride + gb_off + C + c]);
        dgb[(size_t)n * gb_stride + gb_off + c] = __float2bfloat16_rn(s1);
        dgb[(size_t)n * gb_stride + gb_off + C + c] = __float2bfloat16_rn(s2 + sdn * (g3 * s3 + b3 * s1));
      }
      db += sdn * (1.f + gam) * s1;
      dg += sdn * (1.f + gam) * s3;
    }
  }
  sdb[ny][tx] = db;
  sdg[ny][tx] = dg;
  __syncthreads();
  if (ny == 0 && c < C && has_bn) {
    db = dg = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      db += sdb[k][tx];
      dg += sdg[k][tx];
    }
    dgamma3[c] = __float2bfloat16_rn(dg);
    dbeta3[c] = __float2bfloat16_rn(db);
    c1[c] = db * inv_count;
    c2[c] = dg * inv_count;
  }
}

// Pass 3: dY = a3 * (du - c1 - yhat * c2) with du = sd * (1 + gamma) * dOut   (no BN: dY = (1 + gamma) * dOut);
//         dskip = (1 + gamma) * dOut when the block has a residual connection.
__global__ void __launch_bounds__(CNN_THREADS) block_finalize_bwd_apply_kernel(const bf16* __restrict__ dout, const bf16* __restrict__ y,
                                                                               const bf16* __restrict__ gb, long gb_stride, long gb_off,
                                                                               const float* __restrict__ sd, const float* __restrict__ a,
                                                                               const float* __restrict__ mean, const float* __restrict__ rstd,
                                                                               const float* __restrict__ c1, const float* __restrict__ c2,
                                                                               bf16* __restrict__ dy, bf16* __restrict__ dskip, int HW, int C) {
  pdl_enter();
  const ChanMap m = make_chan_map(C / 8);
  if (!m.active) return;
  const int n = blockIdx.y;
  int rb, re;
  cta_rows(HW, rb, re);
  float g1[8], fa[8], mu[8], rs[8], k1[8], k2[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    g1[k] = 1.f; fa[k] = 1.f; mu[k] = 0.f; rs[k] = 1.f; k1[k] = 0.f; k2[k] = 0.f;
  }
  if (gb != nullptr) {
    float t[8];
    load8bf(gb + (size_t)n * gb_stride + gb_off + C + m.cv * 8, t);
#pragma unroll
    for (int k = 0; k < 8; ++k) g1[k] = 1.f + t[k];
  }
  const bool has_bn = a != nullptr;
  if (has_bn) {
    load8f(a + m.cv * 8, fa);
    load8f(mean + m.cv * 8, mu);
    load8f(rstd + m.cv * 8, rs);
    load8f(c1 + m.cv * 8, k1);
    load8f(c2 + m.cv * 8, k2);
  }
  const float sdn = sd != nullptr ? sd[n] : 1.f;
  const size_t off = ((size_t)n * HW) * C + (size_t)m.cv * 8;
  for (int r = rb + m.r0; r < re; r += m.rows_per_pass) {
    float g[8], f[8], o[8];
    load8bf(dout + off + (size_t)r * C, g);
    if (has_bn) load8bf(y + off + (size_t)r * C, f);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const float dz = g1[k] * g[k];
      g[k] = dz;
      o[k] = has_bn ? fa[k] * (sdn * dz - k1[k] - (f[k] - mu[k]) * rs[k] * k2[k]) : dz;
    }
    store8bf(dy + off + (size_t)r * C, o);
    if (dskip != nullptr) store8bf(dskip + off + (size_t)r * C, g);
  }
}

// S: fp32 [3, N, C] scratch, zeroed by this call.
RT1_API int rt1_block_finalize_bwd(const void* dout, const void* y, const void* skip, const void* gb, long long gb_stride,
                                   long long gb_off, const float* sd, const float* a, const float* mean, const float* rstd,
                                   const void* gamma3, const void* beta3, float* S, float* c1, float* c2, void* dgb, void* dgamma3,
                                   void* dbeta3, void* dy, void* dskip, int N, int HW, int C, cudaStream_t stream) {
  RT1_REQUIRE(C % 8 == 0, "rt1_block_finalize_bwd: C=%d must be a multiple of 8", C);
  RT1_CHECK_CUDA(cudaMemsetAsync(S, 0, sizeof(float) * 3 * (size_t)N * C, stream));
  dim3 grid(cnn_ctas_per_frame(N, HW, C), N, cnn_chan_chunks(C));
  rt1_launch(block_finalize_bwd_reduce_kernel, grid, CNN_THREADS, 0, stream, (const bf16*)dout, (const bf16*)y, (const bf16*)skip, mean, rstd, S,
                                                                     HW, C, N);
  int rc = rt1_check_launch("rt1_block_finalize_bwd(reduce)");
  if (rc) return rc;
  rt1_launch(block_finalize_bwd_combine_kernel, rt1_cdiv(C, 32), dim3(32, 8), 0, stream, S, (const bf16*)gb, gb_stride, gb_off, sd, (const bf16*)gamma3,
                                                                         (const bf16*)beta3, (bf16*)dgb, (bf16*)dgamma3, (bf16*)dbeta3, c1,
                                                                         c2, N, C, 1.f / ((float)N * HW));
  rc = rt1_check_launch("rt1_block_finalize_bwd(combine)");
  if (rc) return rc;
  rt1_launch(block_finalize_bwd_apply_kernel, grid, CNN_THREADS, 0, stream, (const bf16*)dout, (const bf16*)y, (const bf16*)gb, gb_stride, gb_off, sd,
                                                                    a, mean, rstd, c1, c2, (bf16*)dy, (bf16*)dskip, HW, C);
  return rt1_check_launch("rt1_block_finalize_bwd(apply)");
}

// ------------------------------------------------------------------------------- SE excite backward: ds[n, c]
// ds_sum[n, c] = sum_hw G * SiLU(a*x + b)      (G = gradient w.r.t. the SE-scaled tensor)
__global__ void __launch_bounds__(CNN_THREADS) se_scale_bwd_reduce_kernel(const bf16* __restrict__ g, const bf16* __restrict__ x,
                                                                          const float* __restrict__ a, const float* __restrict__ b,
                                                                          float* __restrict__ ds, int HW, int C) {
  pdl_enter();
  __shared__ float sm[CNN_THREADS * 8];
  const ChanMap m = make_chan_map(C / 8);
  const int n = blockIdx.y;
  int rb, re;
  cta_rows(HW, rb, re);
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (m.active) {
    float fa[8], fb[8];
    load8f(a + m.cv * 8, fa);
    load8f(b + m.cv * 8, fb);
    const size_t off = ((size_t)n * HW) * C + (size_t)m.cv * 8;
    for (int r = rb + m.r0; r < re; r += m.rows_per_pass) {
      float gg[8], f[8];
      load8bf(g + off + (size_t)r * C, gg);
      load8bf(x + off + (size_t)r * C, f);
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] += gg[k] * silu_f(fmaf(fa[k], f[k], fb[k]));
    }
  }
  cta_channel_reduce_atomic(m, acc, sm, ds + (size_t)n * C);
}

RT1_API int rt1_se_scale_bwd_reduce(const void* g, const void* x, const float* a, const float* b, float* ds, int N, int HW, int C,
                                    cudaStream_t stream) {
  RT1_REQUIRE(C % 8 == 0, "rt1_se_scale_bwd_reduce: C=%d must be a multiple of 8", C);
  RT1_CHECK_CUDA(cudaMemsetAsync(ds, 0, sizeof(float) * (size_t)N * C, stream));
  dim3 grid(cnn_ctas_per_frame(N, HW, C), N, cnn_chan_chunks(C));
  rt1_launch(se_scale_bwd_reduce_kernel, grid, CNN_THREADS, 0, stream, (const bf16*)g, (const bf16*)x, a, b, ds, HW, C);
  return rt1_check_launch("rt1_se_scale_bwd_reduce");
}

// ---- one pass for the SE-scale gradient AND the BatchNorm-backward sums of the same tensor ---------------------------------
// With t = SiLU(u) * s[n,c], u = a x + b, and the pooled-gradient term dpool[n,c]:  dT = G s + dpool,  dh = dT SiLU'(u).
// s and dpool are constant over the frame, so the sums BatchNorm's backward needs separate into per-(frame, channel) parts
//   sum_hw dh       = s * A1 + dpool * A2        A1 = sum G SiLU'(u)        A2 = sum SiLU'(u)
//   sum_hw dh xhat  = s * A3 + dpool * A4        A3 = sum G SiLU'(u) xhat   A4 = sum SiLU'(u) xhat
// which can be taken in the SAME pass over (G, x) that produces ds = sum G SiLU(u) for the squeeze-excite backward —
// before s's gradient path (SE backward → dpool) has run.  This replaces the separate reduce pass of rt1_bn_act_bwd
// (one full read of G and x per block).  out: fp32 [5][N][C] = ds, A1, A2, A3, A4 (zeroed here).
__global__ void __launch_bounds__(CNN_THREADS) se_bn_bwd_reduce_kernel(const bf16* __restrict__ g, const bf16* __restrict__ x,
                                                                       const float* __restrict__ a, const float* __restrict__ b,
                                                                       const float* __restrict__ mean, const float* __restrict__ rstd,
                                                                       float* __restrict__ out, int N, int HW, int C) {
  pdl_enter();
  __shared__ float sm[CNN_THREADS * 8];
  const ChanMap m = make_chan_map(C / 8);
  const int n = blockIdx.y;
  int rb, re;
  cta_rows(HW, rb, re);
  p2 pds[4], pa1[4], pa2[4], pa3[4], pa4[4];          // channel pairs: every operation below is one packed instruction
#pragma unroll
  for (int j = 0; j < 4; ++j) pds[j] = pa1[j] = pa2[j] = pa3[j] = pa4[j] = 0;
  if (m.active) {
    p2 fa[4], fb[4], rs[4], nmr[4];
    {
      p2 mu[4];
      load8p(a + m.cv * 8, fa);
      load8p(b + m.cv * 8, fb);
      load8p(mean + m.cv * 8, mu);
      load8p(rstd + m.cv * 8, rs);
      const p2 neg1 = p2_pack(-1.f, -1.f);
#pragma unroll
      for (int j = 0; j < 4; ++j) nmr[j] = p2_mul(p2_mul(mu[j], rs[j]), neg1);      // xhat = x rstd - mean rstd
    }
    const p2 one = p2_pack(1.f, 1.f), neg1 = p2_pack(-1.f, -1.f);
    const size_t off = ((size_t)n * HW) * C + (size_t)m.cv * 8;
    // four rows per trip: eight 16-byte loads are in flight before the first one is consumed (one row per trip left the
    // kernel latency-bound at 22 % of the copy bandwidth)
    constexpr int RU = 4;
    for (int r = rb + m.r0; r < re; r += RU * m.rows_per_pass) {
      bf16x8 rawg[RU], rawx[RU];
#pragma unroll
      for (int q = 0; q < RU; ++q) {
        const int rq = r + q * m.rows_per_pass;
        if (rq < re) {
          rawg[q] = *reinterpret_cast<const bf16x8*>(g + off + (size_t)rq * C);
          rawx[q] = *reinterpret_cast<const bf16x8*>(x + off + (size_t)rq * C);
        }
      }
#pragma unroll
      for (int q = 0; q < RU; ++q) {
        if (r + q * m.rows_per_pass >= re) continue;
        p2 gg[4], f[4];
        unpack8p(rawg[q], gg);
        unpack8p(rawx[q], f);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const p2 u = p2_fma(fa[j], f[j], fb[j]);
          const p2 sg = p2_sigmoid(u);
          const p2 us = p2_mul(u, sg);
          const p2 d1 = p2_fma(us, p2_fma(sg, neg1, one), sg);      // SiLU'(u) = s + u s (1 - s)
          const p2 xh = p2_fma(f[j], rs[j], nmr[j]);
          const p2 gd = p2_mul(gg[j], d1);
          pds[j] = p2_fma(gg[j], us, pds[j]);
          pa1[j] = p2_add(pa1[j], gd);
          pa2[j] = p2_add(pa2[j], d1);
          pa3[j] = p2_fma(gd, xh, pa3[j]);
          pa4[j] = p2_fma(d1, xh, pa4[j]);
        }
      }
    }
  }
  float ds[8], a1[8], a2[8], a3[8], a4[8];
  p2_to8(pds, ds);
  p2_to8(pa1, a1);
  p2_to8(pa2, a2);
  p2_to8(pa3, a3);
  p2_to8(pa4, a4);
  const size_t plane = (size_t)N * C, row = (size_t)n * C;
  cta_channel_reduce_atomic(m, ds, sm, out + row);
  cta_channel_reduce_atomic(m, a1, sm, out + plane + row);
  cta_channel_reduce_atomic(m, a2, sm, out + 2 * plane + row);
  cta_channel_reduce_atomic(m, a3, sm, out + 3 * plane + row);
  cta_channel_reduce_atomic(m, a4, sm, out + 4 * plane + row);
}

RT1_API int rt1_se_bn_bwd_reduce(const void* g, const void* x, const float* a, const float* b, const float* mean, const float* rstd,
                                 float* out5, int N, int HW, int C, cudaStream_t stream) {
  RT1_REQUIRE(C % 8 == 0, "rt1_se_bn_bwd_reduce: C=%d must be a multiple of 8", C);
  RT1_CHECK_CUDA(cudaMemsetAsync(out5, 0, sizeof(float) * 5 * (size_t)N * C, stream));
  dim3 grid(cnn_ctas_per_frame(N, HW, C), N, cnn_chan_chunks(C));
  rt1_launch(se_bn_bwd_reduce_kernel, grid, CNN_THREADS, 0, stream, (const bf16*)g, (const bf16*)x, a, b, mean, rstd, out5, N, HW, C);
  return rt1_check_launch("rt1_se_bn_bwd_reduce");
}

// ---------------------------------------------------------------- backward of  t = act(BN(x)) [* s + pooled grad]
// dT = G * s[n,c] + dpool[n,c];  dh = silu ? dT * SiLU'(a*x+b) : dT;  xhat = (x - mean) * rstd
// reduce: sums[0, c] += dh, sums[1, c] += dh * xhat
// apply:  dX = a * (dh - c1 - xhat * c2) with c1 = sums[0,c] / M, c2 = sums[1,c] / M taken from the sums by every thread
//         (two loads per channel — a launch of its own for this used to sit between the two passes of all 54 BatchNorm
//         backwards of a step); the CTAs of frame 0 / row range 0 also write dbeta = sums[0], dgamma = sums[1].
template <bool APPLY>
__global__ void __launch_bounds__(CNN_THREADS, APPLY ? 3 : 2) bn_act_bwd_kernel(const bf16* __restrict__ g, const bf16* __restrict__ x,
                                                                 const float* __restrict__ a, const float* __restrict__ b,
                                                                 const float* __restrict__ mean, const float* __restrict__ rstd,
                                                                 const float* __restrict__ s, const float* __restrict__ dpool,
                                                                 float* __restrict__ sums, float inv_count, bf16* __restrict__ dgamma,
                                                                 bf16* __restrict__ dbeta, bf16* __restrict__ dx, int HW, int C, int silu) {
  pdl_enter();
  __shared__ float sm[APPLY ? 1 : CNN_THREADS * 8];
  const ChanMap m = make_chan_map(C / 8);
  const int n = blockIdx.y;
  int rb, re;
  cta_rows(HW, rb, re);
  float acc1[8] = {0, 0, 0, 0, 0, 0, 0, 0}, acc2[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (m.active) {
    float fa[8], fb[8], mu[8], rs[8], fs[8], dp[8], k1[8], k2[8];
    load8f(a + m.cv * 8, fa);
    load8f(b + m.cv * 8, fb);
    load8f(mean + m.cv * 8, mu);
    load8f(rstd + m.cv * 8, rs);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      fs[k] = 1.f; dp[k] = 0.f; k1[k] = 0.f; k2[k] = 0.f;
    }
    if (s != nullptr) load8f(s + (size_t)n * C + m.cv * 8, fs);
    if (dpool != nullptr) load8f(dpool + (size_t)n * C + m.cv * 8, dp);
    if (APPLY) {
      load8f(sums + m.cv * 8, k1);
      load8f(sums + C + m.cv * 8, k2);
      if (blockIdx.x == 0 && n == 0 && m.r0 == 0) {
        store8bf(dbeta + m.cv * 8, k1);
        store8bf(dgamma + m.cv * 8, k2);
      }
      // dX = a (dh - c1 - xhat c2) = a dh + P x + Q with P = -a rstd c2, Q = -a c1 - P mean: two constants per channel
      // instead of four (mean, rstd, c1, c2) keeps the thread under 100 registers
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const float pk = -fa[k] * rs[k] * k2[k] * inv_count;
        k2[k] = -fa[k] * k1[k] * inv_count - pk * mu[k];     // Q
        k1[k] = pk;                                          // P
      }
    }
    const size_t off = ((size_t)n * HW) * C + (size_t)m.cv * 8;
    // two rows per trip: four 16-byte loads are in flight before the first one is consumed
    for (int r = rb + m.r0; r < re; r += 2 * m.rows_per_pass) {
      const int r2 = r + m.rows_per_pass;
      const bool has2 = r2 < re;
      bf16x8 raw[4];
      raw[0] = *reinterpret_cast<const bf16x8*>(g + off + (size_t)r * C);
      raw[1] = *reinterpret_cast<const bf16x8*>(x + off + (size_t)r * C);
      if (has2) {
        raw[2] = *reinterpret_cast<const bf16x8*>(g + off + (size_t)r2 * C);
        raw[3] = *reinterpret_cast<const bf16x8*>(x + off + (size_t)r2 * C);
      }
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        if (h == 1 && !has2) break;
        float gg[8], f[8], o[8];
        unpack8(raw[2 * h], gg);
        unpack8(raw[2 * h + 1], f);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const float dT = fmaf(gg[k], fs[k], dp[k]);
          const float dh = silu ? dT * silu_grad_f(fmaf(fa[k], f[k], fb[k])) : dT;
          if (APPLY) o[k] = fmaf(fa[k], dh, fmaf(k1[k], f[k], k2[k]));
          else {
            acc1[k] += dh;
            acc2[k] += dh * (f[k] - mu[k]) * rs[k];
          }
        }
        if (APPLY) store8bf(dx + off + (size_t)(h ? r2 : r) * C, o);
      }
    }
  }
  if (!APPLY) {
    cta_channel_reduce_atomic(m, acc1, sm, sums);
    cta_channel_reduce_atomic(m, acc2, sm, sums + C);
  }
}

// Full backward of t = act(BN(x)) * s (+ pooled-gradient term): reduce → apply (which also emits dgamma / dbeta).
//   sums: fp32 [2, C].  If `sums_ready` the reduce pass is skipped: the producer — the depthwise backward-data kernel, or
//   rt1_se_bwd_b for a squeeze-excite block — already accumulated the sums for the given `g`.
RT1_API int rt1_bn_act_bwd(const void* g, const void* x, const float* a, const float* b, const float* mean, const float* rstd,
                           const float* s, const float* dpool, float* sums, int sums_ready, void* dgamma, void* dbeta, void* dx,
                           int N, int HW, int C, int silu, cudaStream_t stream) {
  RT1_REQUIRE(C % 8 == 0, "rt1_bn_act_bwd: C=%d must be a multiple of 8", C);
  RT1_REQUIRE((((uintptr_t)dgamma | (uintptr_t)dbeta) & 15) == 0, "rt1_bn_act_bwd: dgamma / dbeta must be 16-byte aligned");
  dim3 grid(cnn_ctas_per_frame(N, HW, C), N, cnn_chan_chunks(C));
  const float inv_count = 1.f / ((float)N * HW);
  if (!sums_ready) {
    RT1_CHECK_CUDA(cudaMemsetAsync(sums, 0, sizeof(float) * 2 * (size_t)C, stream));
    rt1_launch(bn_act_bwd_kernel<false>, grid, CNN_THREADS, 0, stream, (const bf16*)g, (const bf16*)x, a, b, mean, rstd, s, dpool, sums, inv_count,
                                                               nullptr, nullptr, nullptr, HW, C, silu);
    if (int rc = rt1_check_launch("rt1_bn_act_bwd(reduce)")) return rc;
  }
  rt1_launch(bn_act_bwd_kernel<true>, grid, CNN_THREADS, 0, stream, (const bf16*)g, (const bf16*)x, a, b, mean, rstd, s, dpool, sums, inv_count,
                                                            (bf16*)dgamma, (bf16*)dbeta, (bf16*)dx, HW, C, silu);
  return rt1_check_launch("rt1_bn_act_bwd(apply)");
}
