// TokenLearner attention pooling (reference tokenizers/token_learner.py:64-95).
//
// feats [n, P, C] bf16 (P = h*w positions, C = 512)  →  tokens [n, T, C] (T = 8)
//   xn   = LayerNorm_C(feats) ; hid = gelu_tanh(xn W1^T + b1) (64) ; logit = hid W2^T + b2 (T)
//   wmap = softmax over the P positions ; tokens[t] = sum_p wmap[p, t] * feats[p]
// The two products that carry the FLOPs — LN(x)·W1^T [n*P,C]x[C,64] and its two gradients — run on the tcgen05 GEMM over ALL
// frames at once, LayerNorm uses the decoder's kernels, and only the tiny per-frame pieces live here:
//   tl_pool_fwd    : logits = gelu(hpre)·W2^T + b2 → softmax over the P positions → tokens = wmap^T · x
//   tl_pool_bwd    : dwmap = x · dtok^T → softmax backward → dhpre (for the GEMMs), dW2 / db2 partial sums
//   tl_direct_bwd  : dfeats += wmap · dtok   (the pooling path's direct gradient; after LayerNorm backward)
// (Round 1 also had one-CTA-per-frame kernels that did all five products in scalar code on 48 of 148 SMs — 1.8 ms for the
// backward at batch 48 against 0.07 ms for this path; they were removed, the measurement is in profiles/ncu_notes.md.)
#include "common.cuh"

#define TL_HID 64
#define TLP_THREADS 256
#define TLP_MAXP 512   // positions per frame (RT-1 at 256x456: 8 x 15 = 120; 512 covers inputs up to 512 x 1024)
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
  pdl_enter();
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
  pdl_enter();
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
  pdl_enter();
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
  rt1_launch(tl_pool_fwd_kernel, grid, TLP_THREADS, 0, stream, (const bf16*)feats, (const bf16*)hpre, (const bf16*)w2, (const bf16*)b2,
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
  rt1_launch(tl_pool_bwd_kernel, n, 512, smem, stream, (const bf16*)feats, (const bf16*)hpre, wmap, (const bf16*)w2, (const bf16*)dtokens,
                                               dtok_stride, (bf16*)dhpre, scratch, P, C, T);
  return rt1_check_launch("rt1_tl_pool_bwd");
}

RT1_API int rt1_tl_direct_bwd(const float* wmap, const void* dtokens, long long dtok_stride, void* dfeats, int n, int P, int C, int T,
                              cudaStream_t stream) {
  RT1_REQUIRE(C % 64 == 0 && T <= TLP_MAXT && P <= TLP_MAXP, "rt1_tl_direct_bwd: unsupported P=%d C=%d T=%d", P, C, T);
  dim3 grid(n, C / 64);
  rt1_launch(tl_direct_bwd_kernel, grid, TLP_THREADS, 0, stream, wmap, (const bf16*)dtokens, dtok_stride, (bf16*)dfeats, P, C, T);
  return rt1_check_launch("rt1_tl_direct_bwd");
}
