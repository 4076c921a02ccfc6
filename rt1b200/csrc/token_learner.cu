// TokenLearner attention pooling, one CTA per frame (replaces LN + 2 tiny convs + GELU + softmax +
// bmm of K9).
//
// feats [n, P, C] bf16 (P = h*w = 120 positions, C = 512)  →  tokens [n, T, C] (T = 8)
//   xn   = LayerNorm_C(feats) ; hid = gelu_tanh(xn W1^T + b1) (64) ; logit = hid W2^T + b2 (T)
//   wmap = softmax over the P positions ; tokens[t] = sum_p wmap[p, t] * feats[p]
// The frame's feature map (61 KB) and W1 (64 KB) live in shared memory: HBM is read once.  The backward
// recomputes the forward internals from the same tile (nothing but the input is saved) and produces dfeats plus
// per-frame parameter-gradient partials that are atomically folded into an fp32 scratch (a second kernel casts
// them into the bf16 gradient slots).
#include "common.cuh"

#define TL_THREADS 256
#define TL_HID 64

struct TLDims {
  int P, C, T;
  float eps;
};

// scratch layout (fp32): dW1 [64*C] | db1 [64] | dW2 [T*64] | db2 [T] | dgamma [C] | dbeta [C]
__host__ __device__ inline long tl_scratch_floats(int C, int T) { return (long)TL_HID * C + TL_HID + (long)T * TL_HID + T + 2L * C; }

struct TLSmem {
  bf16* x;       // [P][C+8]
  const bf16* w1;  // [64][w1ld]: staged in shared memory by the forward kernel, read from global/L2 by the backward
  int w1ld;
  float* gamma;  // [C]
  float* beta;   // [C]
  float* mean;   // [P]
  float* rstd;   // [P]
  float* hpre;   // [P][65]
  float* wmap;   // [P][T]   logits → softmax weights
};

__device__ __forceinline__ TLSmem tl_carve(uint8_t* smem, const TLDims& d, const bf16* w1_global, bool stage_w1) {
  TLSmem s;
  const int ld = d.C + 8;
  s.x = reinterpret_cast<bf16*>(smem);
  bf16* after_x = s.x + (size_t)d.P * ld;
  s.w1 = stage_w1 ? after_x : w1_global;
  s.w1ld = stage_w1 ? ld : d.C;
  s.gamma = reinterpret_cast<float*>(after_x + (stage_w1 ? (size_t)TL_HID * ld : 0));
  s.beta = s.gamma + d.C;
  s.mean = s.beta + d.C;
  s.rstd = s.mean + d.P;
  s.hpre = s.rstd + d.P;
  s.wmap = s.hpre + (size_t)d.P * (TL_HID + 1);
  return s;
}
static size_t tl_smem_bytes(int P, int C, int T, bool bwd) {
  size_t b = (size_t)(P + (bwd ? 0 : TL_HID)) * (C + 8) * 2 + (size_t)(2 * C + 2 * P) * 4 + (size_t)P * (TL_HID + 1) * 4 +
             (size_t)P * T * 4;
  if (bwd) b += (size_t)T * C * 4 /*dtok*/ + (size_t)P * T * 4 /*dlogit*/ + (size_t)P * (TL_HID + 1) * 4 /*dhpre*/ + 16;
  return b + 16;
}

// Shared forward: fills smem tiles and computes hpre (pre-GELU) and wmap (softmax weights).
__device__ void tl_forward_common(const TLSmem& s, const TLDims& d, const bf16* feats, const bf16* ln_w, const bf16* ln_b,
                                  const bf16* w1, const bf16* b1, const bf16* w2, const bf16* b2) {
  const int P = d.P, C = d.C, T = d.T, ld = C + 8, vpr = C / 8;
  for (int i = threadIdx.x; i < P * vpr; i += blockDim.x) {
    const int r = i / vpr, c = i % vpr;
    *reinterpret_cast<uint4*>(s.x + r * ld + c * 8) = *reinterpret_cast<const uint4*>(feats + (size_t)r * C + c * 8);
  }
  if (s.w1 != w1) {
    bf16* w1s = const_cast<bf16*>(s.w1);
    for (int i = threadIdx.x; i < TL_HID * vpr; i += blockDim.x) {
      const int r = i / vpr, c = i % vpr;
      *reinterpret_cast<uint4*>(w1s + r * ld + c * 8) = *reinterpret_cast<const uint4*>(w1 + (size_t)r * C + c * 8);
    }
  }
  for (int i = threadIdx.x; i < C; i += blockDim.x) {
    s.gamma[i] = __bfloat162float(ln_w[i]);
    s.beta[i] = __bfloat162float(ln_b[i]);
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  for (int p = warp; p < P; p += nw) {
    float sum = 0.f;
    for (int c = lane; c < C; c += 32) sum += __bfloat162float(s.x[p * ld + c]);
    const float mean = warp_sum(sum) / C;
    float q = 0.f;
    for (int c = lane; c < C; c += 32) {
      const float t = __bfloat162float(s.x[p * ld + c]) - mean;
      q += t * t;
    }
    const float rstd = rsqrtf(warp_sum(q) / C + d.eps);
    if (lane == 0) {
      s.mean[p] = mean;
      s.rstd[p] = rstd;
    }
  }
  __syncthreads();
  // hpre[p][u] = <LN(x_p), W1_u> + b1[u]
  for (int t = threadIdx.x; t < P * TL_HID; t += blockDim.x) {
    const int p = t / TL_HID, u = t % TL_HID;
    const float mean = s.mean[p], rstd = s.rstd[p];
    float acc = 0.f;
    for (int c = 0; c < C; c += 8) {
      float fx[8], fw[8];
      unpack8(*reinterpret_cast<const bf16x8*>(s.x + p * ld + c), fx);
      unpack8(*reinterpret_cast<const bf16x8*>(s.w1 + u * s.w1ld + c), fw);
#pragma unroll
      for (int e = 0; e < 8; ++e) acc = fmaf((fx[e] - mean) * rstd * s.gamma[c + e] + s.beta[c + e], fw[e], acc);
    }
    s.hpre[p * (TL_HID + 1) + u] = acc + __bfloat162float(b1[u]);
  }
  __syncthreads();
  for (int t = threadIdx.x; t < P * T; t += blockDim.x) {
    const int p = t / T, k = t % T;
    float acc = __bfloat162float(b2[k]);
    for (int u = 0; u < TL_HID; ++u) acc = fmaf(gelu_tanh_f(s.hpre[p * (TL_HID + 1) + u]), __bfloat162float(w2[k * TL_HID + u]), acc);
    s.wmap[p * T + k] = acc;
  }
  __syncthreads();
  for (int k = warp; k < T; k += nw) {  // softmax over positions
    float mx = -INFINITY;
    for (int p = lane; p < P; p += 32) mx = fmaxf(mx, s.wmap[p * T + k]);
    mx = warp_max(mx);
    float sum = 0.f;
    for (int p = lane; p < P; p += 32) {
      const float e = __expf(s.wmap[p * T + k] - mx);
      s.wmap[p * T + k] = e;
      sum += e;
    }
    const float inv = 1.f / warp_sum(sum);
    for (int p = lane; p < P; p += 32) s.wmap[p * T + k] *= inv;
  }
  __syncthreads();
}

__global__ void __launch_bounds__(TL_THREADS) tl_fwd_kernel(const bf16* __restrict__ feats, const bf16* __restrict__ ln_w,
                                                            const bf16* __restrict__ ln_b, const bf16* __restrict__ w1,
                                                            const bf16* __restrict__ b1, const bf16* __restrict__ w2,
                                                            const bf16* __restrict__ b2, bf16* __restrict__ tokens, long out_stride,
                                                            TLDims d) {
  extern __shared__ __align__(16) uint8_t smem[];
  const TLSmem s = tl_carve(smem, d, w1, true);
  const int n = blockIdx.x, P = d.P, C = d.C, T = d.T, ld = C + 8, vpr = C / 8;
  tl_forward_common(s, d, feats + (size_t)n * P * C, ln_w, ln_b, w1, b1, w2, b2);
  for (int t = threadIdx.x; t < T * vpr; t += blockDim.x) {
    const int k = t / vpr, c = t % vpr;
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int p = 0; p < P; ++p) {
      const float w = s.wmap[p * T + k];
      float fx[8];
      unpack8(*reinterpret_cast<const bf16x8*>(s.x + p * ld + c * 8), fx);
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] = fmaf(w, fx[e], acc[e]);
    }
    *reinterpret_cast<bf16x8*>(tokens + (size_t)n * out_stride + (size_t)k * C + c * 8) = pack8(acc);
  }
}

__global__ void __launch_bounds__(TL_THREADS) tl_bwd_kernel(const bf16* __restrict__ feats, const bf16* __restrict__ ln_w,
                                                            const bf16* __restrict__ ln_b, const bf16* __restrict__ w1,
                                                            const bf16* __restrict__ b1, const bf16* __restrict__ w2,
                                                            const bf16* __restrict__ b2, const bf16* __restrict__ dtokens,
                                                            long dtok_stride, bf16* __restrict__ dfeats, float* __restrict__ scratch,
                                                            TLDims d) {
  extern __shared__ __align__(16) uint8_t smem[];
  const TLSmem s = tl_carve(smem, d, w1, false);
  const int n = blockIdx.x, P = d.P, C = d.C, T = d.T, ld = C + 8, vpr = C / 8, HL = TL_HID + 1;
  float* dtok = s.wmap + (size_t)P * T;      // [T][C]
  float* dlogit = dtok + (size_t)T * C;      // [P][T]
  float* dhpre = dlogit + (size_t)P * T;     // [P][65]
  tl_forward_common(s, d, feats + (size_t)n * P * C, ln_w, ln_b, w1, b1, w2, b2);
  for (int i = threadIdx.x; i < T * C; i += blockDim.x)
    dtok[i] = __bfloat162float(dtokens[(size_t)n * dtok_stride + i]);
  __syncthreads();
  // dwmap[p][k] = <dtok_k, x_p>
  for (int t = threadIdx.x; t < P * T; t += blockDim.x) {
    const int p = t / T, k = t % T;
    float acc = 0.f;
    for (int c = 0; c < C; c += 8) {
      float fx[8];
      unpack8(*reinterpret_cast<const bf16x8*>(s.x + p * ld + c), fx);
#pragma unroll
      for (int e = 0; e < 8; ++e) acc = fmaf(fx[e], dtok[k * C + c + e], acc);
    }
    dlogit[p * T + k] = acc;
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  for (int k = warp; k < T; k += nw) {  // softmax backward over positions
    float dot = 0.f;
    for (int p = lane; p < P; p += 32) dot += dlogit[p * T + k] * s.wmap[p * T + k];
    dot = warp_sum(dot);
    for (int p = lane; p < P; p += 32) dlogit[p * T + k] = s.wmap[p * T + k] * (dlogit[p * T + k] - dot);
  }
  __syncthreads();
  float* g_dw1 = scratch;
  float* g_db1 = g_dw1 + (size_t)TL_HID * C;
  float* g_dw2 = g_db1 + TL_HID;
  float* g_db2 = g_dw2 + (size_t)T * TL_HID;
  float* g_dgamma = g_db2 + T;
  float* g_dbeta = g_dgamma + C;
  // dW2[k][u], db2[k], dhpre[p][u]
  for (int t = threadIdx.x; t < T * TL_HID; t += blockDim.x) {
    const int k = t / TL_HID, u = t % TL_HID;
    float acc = 0.f;
    for (int p = 0; p < P; ++p) acc = fmaf(dlogit[p * T + k], gelu_tanh_f(s.hpre[p * HL + u]), acc);
    atomicAdd(g_dw2 + t, acc);
  }
  for (int k = threadIdx.x; k < T; k += blockDim.x) {
    float acc = 0.f;
    for (int p = 0; p < P; ++p) acc += dlogit[p * T + k];
    atomicAdd(g_db2 + k, acc);
  }
  for (int t = threadIdx.x; t < P * TL_HID; t += blockDim.x) {
    const int p = t / TL_HID, u = t % TL_HID;
    float acc = 0.f;
    for (int k = 0; k < T; ++k) acc = fmaf(dlogit[p * T + k], __bfloat162float(w2[k * TL_HID + u]), acc);
    dhpre[p * HL + u] = acc * gelu_tanh_grad_f(s.hpre[p * HL + u]);
  }
  __syncthreads();
  for (int u = threadIdx.x; u < TL_HID; u += blockDim.x) {
    float acc = 0.f;
    for (int p = 0; p < P; ++p) acc += dhpre[p * HL + u];
    atomicAdd(g_db1 + u, acc);
  }
  // dW1[u][c..c+8] = sum_p dhpre[p][u] * xn[p][c..]
  for (int t = threadIdx.x; t < TL_HID * vpr; t += blockDim.x) {
    const int u = t / vpr, c = (t % vpr) * 8;
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int p = 0; p < P; ++p) {
      const float g = dhpre[p * HL + u], mean = s.mean[p], rstd = s.rstd[p];
      float fx[8];
      unpack8(*reinterpret_cast<const bf16x8*>(s.x + p * ld + c), fx);
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] = fmaf(g, (fx[e] - mean) * rstd * s.gamma[c + e] + s.beta[c + e], acc[e]);
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) atomicAdd(g_dw1 + (size_t)u * C + c + e, acc[e]);
  }
  // dxn[p][c] = sum_u dhpre[p][u] W1[u][c]; LayerNorm backward row-wise; add the direct pooling path.
  // One warp per position; each lane owns columns lane*8 + 256*i.
  float dgam[16], dbet[16];  // C <= 512: two 8-wide chunks per lane
#pragma unroll
  for (int i = 0; i < 16; ++i) dgam[i] = dbet[i] = 0.f;
  for (int p = warp; p < P; p += nw) {
    const float mean = s.mean[p], rstd = s.rstd[p];
    float dxn[16], xh[16];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int c = lane * 8 + 256 * i;
      if (c < C) {
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int u = 0; u < TL_HID; ++u) {
          const float g = dhpre[p * HL + u];
          float fw[8];
          unpack8(*reinterpret_cast<const bf16x8*>(s.w1 + u * s.w1ld + c), fw);
#pragma unroll
          for (int e = 0; e < 8; ++e) acc[e] = fmaf(g, fw[e], acc[e]);
        }
        float fx[8];
        unpack8(*reinterpret_cast<const bf16x8*>(s.x + p * ld + c), fx);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float h = (fx[e] - mean) * rstd;
          xh[8 * i + e] = h;
          dgam[8 * i + e] += acc[e] * h;
          dbet[8 * i + e] += acc[e];
          const float gg = acc[e] * s.gamma[c + e];
          dxn[8 * i + e] = gg;
          s1 += gg;
          s2 += gg * h;
        }
      }
    }
    s1 = warp_sum(s1) / C;
    s2 = warp_sum(s2) / C;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int c = lane * 8 + 256 * i;
      if (c < C) {
        float o[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          float direct = 0.f;
          for (int k = 0; k < T; ++k) direct = fmaf(s.wmap[p * T + k], dtok[k * C + c + e], direct);
          o[e] = rstd * (dxn[8 * i + e] - s1 - xh[8 * i + e] * s2) + direct;
        }
        *reinterpret_cast<bf16x8*>(dfeats + ((size_t)n * P + p) * C + c) = pack8(o);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int c = lane * 8 + 256 * i;
    if (c < C)
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        atomicAdd(g_dgamma + c + e, dgam[8 * i + e]);
        atomicAdd(g_dbeta + c + e, dbet[8 * i + e]);
      }
  }
}

__global__ void tl_param_grad_kernel(const float* __restrict__ scratch, int C, int T, bf16* dw1, bf16* db1, bf16* dw2, bf16* db2,
                                     bf16* dgamma, bf16* dbeta) {
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const long n1 = (long)TL_HID * C, n2 = n1 + TL_HID, n3 = n2 + (long)T * TL_HID, n4 = n3 + T, n5 = n4 + C, n6 = n5 + C;
  if (i >= n6) return;
  const bf16 v = __float2bfloat16_rn(scratch[i]);
  if (i < n1) dw1[i] = v;
  else if (i < n2) db1[i - n1] = v;
  else if (i < n3) dw2[i - n2] = v;
  else if (i < n4) db2[i - n3] = v;
  else if (i < n5) dgamma[i - n4] = v;
  else dbeta[i - n5] = v;
}

RT1_API long long rt1_token_learner_scratch_floats(int C, int T) { return tl_scratch_floats(C, T); }

RT1_API int rt1_token_learner_fwd(const void* feats, const void* ln_w, const void* ln_b, const void* w1, const void* b1,
                                  const void* w2, const void* b2, void* tokens, long long out_stride, int n, int P, int C, int T,
                                  float eps, cudaStream_t stream) {
  RT1_REQUIRE(C % 8 == 0 && C <= 512 && T <= 16, "rt1_token_learner_fwd: unsupported C=%d T=%d", C, T);
  const size_t smem = tl_smem_bytes(P, C, T, false);
  RT1_REQUIRE(smem <= 227 * 1024, "rt1_token_learner_fwd: P=%d C=%d needs %zu B of shared memory", P, C, smem);
  if (int rc = rt1_ensure_dyn_smem((const void*)tl_fwd_kernel, smem)) return rc;
  TLDims d{P, C, T, eps};
  tl_fwd_kernel<<<n, TL_THREADS, smem, stream>>>((const bf16*)feats, (const bf16*)ln_w, (const bf16*)ln_b, (const bf16*)w1,
                                                 (const bf16*)b1, (const bf16*)w2, (const bf16*)b2, (bf16*)tokens, out_stride, d);
  return rt1_check_launch("rt1_token_learner_fwd");
}

// scratch: fp32 [rt1_token_learner_scratch_floats(C, T)], zeroed by this call
RT1_API int rt1_token_learner_bwd(const void* feats, const void* ln_w, const void* ln_b, const void* w1, const void* b1,
                                  const void* w2, const void* b2, const void* dtokens, long long dtok_stride, void* dfeats,
                                  float* scratch, void* dln_w, void* dln_b, void* dw1, void* db1, void* dw2, void* db2, int n, int P,
                                  int C, int T, float eps, cudaStream_t stream) {
  RT1_REQUIRE(C % 8 == 0 && C <= 512 && T <= 16, "rt1_token_learner_bwd: unsupported C=%d T=%d", C, T);
  const size_t smem = tl_smem_bytes(P, C, T, true);
  RT1_REQUIRE(smem <= 227 * 1024, "rt1_token_learner_bwd: P=%d C=%d needs %zu B of shared memory", P, C, smem);
  if (int rc = rt1_ensure_dyn_smem((const void*)tl_bwd_kernel, smem)) return rc;
  const long ns = tl_scratch_floats(C, T);
  RT1_CHECK_CUDA(cudaMemsetAsync(scratch, 0, ns * sizeof(float), stream));
  TLDims d{P, C, T, eps};
  tl_bwd_kernel<<<n, TL_THREADS, smem, stream>>>((const bf16*)feats, (const bf16*)ln_w, (const bf16*)ln_b, (const bf16*)w1,
                                                 (const bf16*)b1, (const bf16*)w2, (const bf16*)b2, (const bf16*)dtokens, dtok_stride,
                                                 (bf16*)dfeats, scratch, d);
  int rc = rt1_check_launch("rt1_token_learner_bwd");
  if (rc) return rc;
  tl_param_grad_kernel<<<rt1_cdiv(ns, 256), 256, 0, stream>>>(scratch, C, T, (bf16*)dw1, (bf16*)db1, (bf16*)dw2, (bf16*)db2,
                                                              (bf16*)dln_w, (bf16*)dln_b);
  return rt1_check_launch("rt1_token_learner_param_grad");
}

// =================================================================================================================
// GEMM-composed TokenLearner (the default path).  The one-CTA-per-frame kernels above do five small matrix products
// in scalar code on 48 of 148 SMs (1.8 ms for the backward at batch 48).  Here the two products that carry the FLOPs
// — LN(x)·W1^T [n*P,C]x[C,64] and its two gradients — run on the tcgen05 GEMM over ALL frames at once, LayerNorm uses
// the decoder's kernels, and only the tiny per-frame pieces stay here:
//   tl_pool_fwd    : logits = gelu(hpre)·W2^T + b2 → softmax over the P positions → tokens = wmap^T · x
//   tl_pool_bwd    : dwmap = x · dtok^T → softmax backward → dhpre (for the GEMMs), dW2 / db2 partial sums
//   tl_direct_bwd  : dfeats += wmap · dtok   (the pooling path's direct gradient; after LayerNorm backward)
#define TLP_THREADS 256
#define TLP_MAXP 128   // positions per frame (RT-1: 8 x 15 = 120)
#define TLP_MAXT 8

// logits → softmax weights for one frame, result in `wm` [P][T] (shared); all threads must call
__device__ void tlp_softmax_weights(const bf16* __restrict__ hpre, const bf16* __restrict__ w2, const bf16* __restrict__ b2,
                                    float* wm, float* w2s, int P, int T) {
  for (int i = threadIdx.x; i < T * TL_HID; i += blockDim.x) w2s[i] = __bfloat162float(w2[i]);
  __syncthreads();
  for (int p = threadIdx.x; p < P; p += blockDim.x) {
    float acc[TLP_MAXT];
#pragma unroll
    for (int k = 0; k < TLP_MAXT; ++k) acc[k] = k < T ? __bfloat162float(b2[k]) : 0.f;
    for (int u0 = 0; u0 < TL_HID; u0 += 8) {
      float h[8];
      unpack8(*reinterpret_cast<const bf16x8*>(hpre + (size_t)p * TL_HID + u0), h);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float gl = gelu_tanh_f(h[e]);
#pragma unroll
        for (int k = 0; k < TLP_MAXT; ++k)
          if (k < T) acc[k] = fmaf(gl, w2s[k * TL_HID + u0 + e], acc[k]);
      }
    }
#pragma unroll
    for (int k = 0; k < TLP_MAXT; ++k)
      if (k < T) wm[p * T + k] = acc[k];
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  for (int k = warp; k < T; k += nw) {
    float mx = -INFINITY;
    for (int p = lane; p < P; p += 32) mx = fmaxf(mx, wm[p * T + k]);
    mx = warp_max(mx);
    float sum = 0.f;
    for (int p = lane; p < P; p += 32) {
      const float e = __expf(wm[p * T + k] - mx);
      wm[p * T + k] = e;
      sum += e;
    }
    const float inv = 1.f / warp_sum(sum);
    for (int p = lane; p < P; p += 32) wm[p * T + k] *= inv;
  }
  __syncthreads();
}

// grid (n, C / 64): every CTA recomputes the frame's softmax weights (61 K MACs) and pools a 64-channel slice
__global__ void __launch_bounds__(TLP_THREADS) tl_pool_fwd_kernel(const bf16* __restrict__ feats, const bf16* __restrict__ hpre,
                                                                  const bf16* __restrict__ w2, const bf16* __restrict__ b2,
                                                                  bf16* __restrict__ tokens, long out_stride, float* __restrict__ wmap,
                                                                  int P, int C, int T) {
  __shared__ float wm[TLP_MAXP * TLP_MAXT];
  __shared__ float w2s[TLP_MAXT * TL_HID];
  __shared__ float part[4][TLP_MAXT * 8][8];
  const int n = blockIdx.x, c0 = blockIdx.y * 64;
  tlp_softmax_weights(hpre + (size_t)n * P * TL_HID, w2, b2, wm, w2s, P, T);
  if (blockIdx.y == 0)
    for (int i = threadIdx.x; i < P * T; i += blockDim.x) wmap[(size_t)n * P * T + i] = wm[i];
  // thread ↔ (quarter of the positions, token k, channel vector): 4 x 8 x 8 = 256
  const int q = threadIdx.x >> 6, k = (threadIdx.x >> 3) & 7, cv = threadIdx.x & 7;
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (k < T) {
    const bf16* xf = feats + (size_t)n * P * C + c0 + cv * 8;
    for (int p = q; p < P; p += 4) {
      const float w = wm[p * T + k];
      float fx[8];
      unpack8(*reinterpret_cast<const bf16x8*>(xf + (size_t)p * C), fx);
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] = fmaf(w, fx[e], acc[e]);
    }
  }
#pragma unroll
  for (int e = 0; e < 8; ++e) part[q][k * 8 + cv][e] = acc[e];
  __syncthreads();
  if (q == 0 && k < T) {
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = part[0][k * 8 + cv][e] + part[1][k * 8 + cv][e] + part[2][k * 8 + cv][e] + part[3][k * 8 + cv][e];
    *reinterpret_cast<bf16x8*>(tokens + (size_t)n * out_stride + (size_t)k * C + c0 + cv * 8) = pack8(acc);
  }
}

// grid (n): dwmap → softmax backward → dhpre [n*P, 64] bf16, dW2 / db2 sums into scratch [T*64 + T] (atomics)
__global__ void __launch_bounds__(512) tl_pool_bwd_kernel(const bf16* __restrict__ feats, const bf16* __restrict__ hpre,
                                                          const float* __restrict__ wmap, const bf16* __restrict__ w2,
                                                          const bf16* __restrict__ dtokens, long dtok_stride, bf16* __restrict__ dhpre,
                                                          float* __restrict__ scratch, int P, int C, int T) {
  extern __shared__ __align__(16) uint8_t tl_dyn[];
  bf16* dts = reinterpret_cast<bf16*>(tl_dyn);                       // [T][C] bf16
  float* dl = reinterpret_cast<float*>(dts + (size_t)T * C);          // [P][T] dwmap → dlogit
  float* wm = dl + TLP_MAXP * TLP_MAXT;                               // [P][T]
  float* w2s = wm + TLP_MAXP * TLP_MAXT;                              // [T][64]
  const int n = blockIdx.x;
  for (int i = threadIdx.x; i < T * C / 8; i += blockDim.x)
    reinterpret_cast<uint4*>(dts)[i] = *reinterpret_cast<const uint4*>(dtokens + (size_t)n * dtok_stride + (size_t)i * 8);
  for (int i = threadIdx.x; i < P * T; i += blockDim.x) wm[i] = wmap[(size_t)n * P * T + i];
  for (int i = threadIdx.x; i < T * TL_HID; i += blockDim.x) w2s[i] = __bfloat162float(w2[i]);
  __syncthreads();
  // dwmap[p][k] = <x_p, dtok_k>: a warp per (p, k) pair, lanes over channels
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  for (int p = warp; p < P; p += nw) {
    const bf16* xr = feats + ((size_t)n * P + p) * C;
    float acc[TLP_MAXT];
#pragma unroll
    for (int k = 0; k < TLP_MAXT; ++k) acc[k] = 0.f;
    for (int c = lane * 8; c < C; c += 256) {
      float fx[8];
      unpack8(*reinterpret_cast<const bf16x8*>(xr + c), fx);
#pragma unroll
      for (int k = 0; k < TLP_MAXT; ++k) {
        if (k < T) {
          float fd[8];
          unpack8(*reinterpret_cast<const bf16x8*>(dts + (size_t)k * C + c), fd);
#pragma unroll
          for (int e = 0; e < 8; ++e) acc[k] = fmaf(fx[e], fd[e], acc[k]);
        }
      }
    }
#pragma unroll
    for (int k = 0; k < TLP_MAXT; ++k) {
      const float s = warp_sum(acc[k]);
      if (lane == 0 && k < T) dl[p * T + k] = s;
    }
  }
  __syncthreads();
  for (int k = warp; k < T; k += nw) {  // softmax backward over positions
    float dot = 0.f;
    for (int p = lane; p < P; p += 32) dot += dl[p * T + k] * wm[p * T + k];
    dot = warp_sum(dot);
    for (int p = lane; p < P; p += 32) dl[p * T + k] = wm[p * T + k] * (dl[p * T + k] - dot);
  }
  __syncthreads();
  // dhpre[p][u] = gelu'(hpre) * sum_k dlogit[p][k] W2[k][u];   dW2[k][u] = sum_p dlogit[p][k] gelu(hpre[p][u])
  // thread ↔ (u, position phase): u = tid % 64, phase = tid / 64 (8 phases)
  {
    const int u = threadIdx.x & 63, ph = threadIdx.x >> 6, nph = blockDim.x >> 6;
    float gw2[TLP_MAXT];
#pragma unroll
    for (int k = 0; k < TLP_MAXT; ++k) gw2[k] = 0.f;
    for (int p = ph; p < P; p += nph) {
      const size_t idx = ((size_t)n * P + p) * TL_HID + u;
      const float h = __bfloat162float(hpre[idx]);
      const float gl = gelu_tanh_f(h);
      float d = 0.f;
#pragma unroll
      for (int k = 0; k < TLP_MAXT; ++k) {
        if (k < T) {
          const float dlk = dl[p * T + k];
          d = fmaf(dlk, w2s[k * TL_HID + u], d);
          gw2[k] = fmaf(dlk, gl, gw2[k]);
        }
      }
      dhpre[idx] = __float2bfloat16_rn(d * gelu_tanh_grad_f(h));
    }
#pragma unroll
    for (int k = 0; k < TLP_MAXT; ++k)
      if (k < T) atomicAdd(scratch + k * TL_HID + u, gw2[k]);
  }
  for (int k = threadIdx.x; k < T; k += blockDim.x) {
    float s = 0.f;
    for (int p = 0; p < P; ++p) s += dl[p * T + k];
    atomicAdd(scratch + T * TL_HID + k, s);
  }
}

// grid (n, C / 64): dfeats[p][c] += sum_k wmap[p][k] * dtok[k][c]
__global__ void __launch_bounds__(TLP_THREADS) tl_direct_bwd_kernel(const float* __restrict__ wmap, const bf16* __restrict__ dtokens,
                                                                    long dtok_stride, bf16* __restrict__ dfeats, int P, int C, int T) {
  __shared__ float dts[TLP_MAXT][64];
  __shared__ float wm[TLP_MAXP * TLP_MAXT];
  const int n = blockIdx.x, c0 = blockIdx.y * 64;
  for (int i = threadIdx.x; i < T * 64; i += blockDim.x)
    dts[i >> 6][i & 63] = __bfloat162float(dtokens[(size_t)n * dtok_stride + (size_t)(i >> 6) * C + c0 + (i & 63)]);
  for (int i = threadIdx.x; i < P * T; i += blockDim.x) wm[i] = wmap[(size_t)n * P * T + i];
  __syncthreads();
  const int cv = threadIdx.x & 7;
  for (int p = threadIdx.x >> 3; p < P; p += TLP_THREADS / 8) {
    bf16* d = dfeats + ((size_t)n * P + p) * C + c0 + cv * 8;
    float o[8];
    unpack8(*reinterpret_cast<const bf16x8*>(d), o);
#pragma unroll
    for (int k = 0; k < TLP_MAXT; ++k) {
      if (k < T) {
        const float w = wm[p * T + k];
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = fmaf(w, dts[k][cv * 8 + e], o[e]);
      }
    }
    *reinterpret_cast<bf16x8*>(d) = pack8(o);
  }
}

// feats [n,P,C], hpre [n*P,64] (= LN(feats)·W1^T + b1, bf16) → tokens (row n at tokens + n*out_stride), wmap fp32 [n,P,T]
RT1_API int rt1_tl_pool_fwd(const void* feats, const void* hpre, const void* w2, const void* b2, void* tokens, long long out_stride,
                            float* wmap, int n, int P, int C, int T, cudaStream_t stream) {
  RT1_REQUIRE(C % 64 == 0 && T <= TLP_MAXT && P <= TLP_MAXP, "rt1_tl_pool_fwd: unsupported P=%d C=%d T=%d", P, C, T);
  dim3 grid(n, C / 64);
  tl_pool_fwd_kernel<<<grid, TLP_THREADS, 0, stream>>>((const bf16*)feats, (const bf16*)hpre, (const bf16*)w2, (const bf16*)b2,
                                                       (bf16*)tokens, out_stride, wmap, P, C, T);
  return rt1_check_launch("rt1_tl_pool_fwd");
}

// → dhpre [n*P,64] bf16; scratch fp32 [T*64 + T] (zeroed here) receives dW2 and db2
RT1_API int rt1_tl_pool_bwd(const void* feats, const void* hpre, const float* wmap, const void* w2, const void* dtokens,
                            long long dtok_stride, void* dhpre, float* scratch, int n, int P, int C, int T, cudaStream_t stream) {
  RT1_REQUIRE(C % 8 == 0 && T <= TLP_MAXT && P <= TLP_MAXP, "rt1_tl_pool_bwd: unsupported P=%d C=%d T=%d", P, C, T);
  const size_t smem = (size_t)T * C * 2 + (size_t)(2 * TLP_MAXP * TLP_MAXT + TLP_MAXT * TL_HID) * 4;
  RT1_REQUIRE(smem <= 48 * 1024, "rt1_tl_pool_bwd: C=%d needs %zu B of shared memory", C, smem);
  RT1_CHECK_CUDA(cudaMemsetAsync(scratch, 0, sizeof(float) * (size_t)(T * TL_HID + T), stream));
  tl_pool_bwd_kernel<<<n, 512, smem, stream>>>((const bf16*)feats, (const bf16*)hpre, wmap, (const bf16*)w2, (const bf16*)dtokens,
                                               dtok_stride, (bf16*)dhpre, scratch, P, C, T);
  return rt1_check_launch("rt1_tl_pool_bwd");
}

RT1_API int rt1_tl_direct_bwd(const float* wmap, const void* dtokens, long long dtok_stride, void* dfeats, int n, int P, int C, int T,
                              cudaStream_t stream) {
  RT1_REQUIRE(C % 64 == 0 && T <= TLP_MAXT && P <= TLP_MAXP, "rt1_tl_direct_bwd: unsupported P=%d C=%d T=%d", P, C, T);
  dim3 grid(n, C / 64);
  tl_direct_bwd_kernel<<<grid, TLP_THREADS, 0, stream>>>(wmap, (const bf16*)dtokens, dtok_stride, (bf16*)dfeats, P, C, T);
  return rt1_check_launch("rt1_tl_direct_bwd");
}
