// Squeeze-and-excitation fully-connected pair on per-frame vectors (the FC part of K5).
//
//   pooled = pooled_sum / HW ; z1 = W1 pooled + b1 ; hid = SiLU(z1) ; z2 = W2 hid + b2 ; s = sigmoid(z2)
// W1 [Cs, C], W2 [C, Cs] are the 1x1-conv weights of the reference's SeModule (film_efficientnet_encoder.py:142-161).
// One CTA per frame; the vectors are tiny (C <= 2304, Cs <= 96) so everything stays in shared memory.
#include "common.cuh"

#define SE_THREADS 1024
#define SE_MAX_CS 128

// One CTA of 1024 threads per frame.  Every matrix-vector product is laid out so that a warp instruction reads
// CONTIGUOUS weights (ncu on the first version: 12 % of the SMs busy, 2-byte loads with a stride of Cs or C elements):
//   z1[u]   = W1[u,:] · pooled    warp per row u, lanes take 8 consecutive channels (16-byte loads)
//   s[c]    = W2[c,:] · hid       thread per row c, bf16x2 loads along the row (rows of neighbouring threads are adjacent)
__global__ void __launch_bounds__(SE_THREADS) se_fc_fwd_kernel(const float* __restrict__ pooled_sum, float inv_hw,
                                                               const bf16* __restrict__ w1, const bf16* __restrict__ b1,
                                                               const bf16* __restrict__ w2, const bf16* __restrict__ b2,
                                                               float* __restrict__ z1_out, float* __restrict__ s_out, int C, int Cs) {
  extern __shared__ float sm[];  // pooled [C] | hid [Cs]
  float* pooled = sm;
  float* hid = sm + C;
  const int n = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  for (int c = threadIdx.x; c < C; c += blockDim.x) pooled[c] = pooled_sum[(size_t)n * C + c] * inv_hw;
  __syncthreads();
  for (int u = warp; u < Cs; u += nw) {
    float acc = 0.f;
    const bf16* wr = w1 + (size_t)u * C;
    for (int c = lane * 8; c < C; c += 256) {
      float fw[8];
      unpack8(*reinterpret_cast<const bf16x8*>(wr + c), fw);
#pragma unroll
      for (int e = 0; e < 8; ++e) acc = fmaf(fw[e], pooled[c + e], acc);
    }
    acc = warp_sum(acc);
    if (lane == 0) {
      const float z = acc + __bfloat162float(b1[u]);
      z1_out[(size_t)n * Cs + u] = z;
      hid[u] = silu_f(z);
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float acc = __bfloat162float(b2[c]);
    const bf162* wr = reinterpret_cast<const bf162*>(w2 + (size_t)c * Cs);   // Cs is even: rows are 4-byte aligned
    for (int u = 0; u < Cs; u += 2) {
      const float2 w = __bfloat1622float2(wr[u >> 1]);
      acc = fmaf(w.x, hid[u], acc);
      acc = fmaf(w.y, hid[u + 1], acc);
    }
    s_out[(size_t)n * C + c] = sigmoid_f(acc);
  }
}

// ds [N, C] (gradient w.r.t. s) → dz2 [N, C], dz1 [N, Cs], dpooled [N, C] (already divided by HW, i.e. the
// gradient each spatial position of the pooled tensor receives)
//   dhid[u]    = sum_c dz2[c] W2[c,u]   a warp walks rows c, lanes along u (contiguous), partial sums meet in shared memory
//   dpooled[c] = sum_u dz1[u] W1[u,c]   thread per column c: neighbouring threads read neighbouring elements of a row
__global__ void __launch_bounds__(SE_THREADS) se_fc_bwd_kernel(const float* __restrict__ ds, const float* __restrict__ s,
                                                               const float* __restrict__ z1, const bf16* __restrict__ w1,
                                                               const bf16* __restrict__ w2, float inv_hw, float* __restrict__ dz2_out,
                                                               float* __restrict__ dz1_out, float* __restrict__ dpooled, int C, int Cs) {
  extern __shared__ float sm[];  // dz2 [C] | dz1 [Cs] | dhid [Cs]
  float* dz2 = sm;
  float* dz1 = sm + C;
  float* dhid = dz1 + Cs;
  const int n = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const float sv = s[(size_t)n * C + c];
    const float g = ds[(size_t)n * C + c] * sv * (1.f - sv);
    dz2[c] = g;
    dz2_out[(size_t)n * C + c] = g;
  }
  for (int u = threadIdx.x; u < Cs; u += blockDim.x) dhid[u] = 0.f;
  __syncthreads();
  {
    float acc[SE_MAX_CS / 32] = {0.f, 0.f, 0.f, 0.f};
    for (int c = warp; c < C; c += nw) {
      const float g = dz2[c];
      const bf16* wr = w2 + (size_t)c * Cs;
#pragma unroll
      for (int j = 0; j < SE_MAX_CS / 32; ++j) {
        const int u = lane + 32 * j;
        if (u < Cs) acc[j] = fmaf(g, __bfloat162float(wr[u]), acc[j]);
      }
    }
#pragma unroll
    for (int j = 0; j < SE_MAX_CS / 32; ++j) {
      const int u = lane + 32 * j;
      if (u < Cs) atomicAdd(dhid + u, acc[j]);
    }
  }
  __syncthreads();
  for (int u = threadIdx.x; u < Cs; u += blockDim.x) {
    const float g = dhid[u] * silu_grad_f(z1[(size_t)n * Cs + u]);
    dz1[u] = g;
    dz1_out[(size_t)n * Cs + u] = g;
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {  // dpooled[c] = sum_u dz1[u] * W1[u, c]
    float acc = 0.f;
    for (int u = 0; u < Cs; ++u) acc = fmaf(dz1[u], __bfloat162float(w1[(size_t)u * C + c]), acc);
    dpooled[(size_t)n * C + c] = acc * inv_hw;
  }
}

// Parameter gradients, summed over the N frames, written as bf16 (no atomics: one thread per output element).
//   dW2[c,u] = sum_n dz2[n,c] hid[n,u]   db2[c] = sum_n dz2[n,c]   dW1[u,c] = sum_n dz1[n,u] pooled[n,c]   db1[u] = sum_n dz1[n,u]
__global__ void se_wgrad_kernel(const float* __restrict__ dz2, const float* __restrict__ dz1, const float* __restrict__ z1,
                                const float* __restrict__ pooled_sum, float inv_hw, bf16* __restrict__ dw1, bf16* __restrict__ db1,
                                bf16* __restrict__ dw2, bf16* __restrict__ db2, int N, int C, int Cs) {
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const long n_w = (long)C * Cs;
  if (i < n_w) {
    {  // dW2 index i = c*Cs + u
      const int c = (int)(i / Cs), u = (int)(i % Cs);
      float acc = 0.f;
      for (int n = 0; n < N; ++n) acc = fmaf(dz2[(size_t)n * C + c], silu_f(z1[(size_t)n * Cs + u]), acc);
      dw2[i] = __float2bfloat16_rn(acc);
    }
    {  // dW1 index i = u*C + c
      const int u = (int)(i / C), c = (int)(i % C);
      float acc = 0.f;
      for (int n = 0; n < N; ++n) acc = fmaf(dz1[(size_t)n * Cs + u], pooled_sum[(size_t)n * C + c] * inv_hw, acc);
      dw1[i] = __float2bfloat16_rn(acc);
    }
  } else if (i < n_w + C) {
    const int c = (int)(i - n_w);
    float acc = 0.f;
    for (int n = 0; n < N; ++n) acc += dz2[(size_t)n * C + c];
    db2[c] = __float2bfloat16_rn(acc);
  } else if (i < n_w + C + Cs) {
    const int u = (int)(i - n_w - C);
    float acc = 0.f;
    for (int n = 0; n < N; ++n) acc += dz1[(size_t)n * Cs + u];
    db1[u] = __float2bfloat16_rn(acc);
  }
}

RT1_API int rt1_se_fc_fwd(const float* pooled_sum, float inv_hw, const void* w1, const void* b1, const void* w2, const void* b2,
                          float* z1, float* s, int N, int C, int Cs, cudaStream_t stream) {
  RT1_REQUIRE(Cs <= SE_MAX_CS && Cs % 2 == 0 && C % 8 == 0 && C <= 8192, "rt1_se_fc_fwd: unsupported C=%d Cs=%d", C, Cs);
  se_fc_fwd_kernel<<<N, SE_THREADS, (C + Cs) * sizeof(float), stream>>>(pooled_sum, inv_hw, (const bf16*)w1, (const bf16*)b1,
                                                                       (const bf16*)w2, (const bf16*)b2, z1, s, C, Cs);
  return rt1_check_launch("rt1_se_fc_fwd");
}

RT1_API int rt1_se_fc_bwd(const float* ds, const float* s, const float* z1, const float* pooled_sum, float inv_hw, const void* w1,
                          const void* w2, float* dz2, float* dz1, float* dpooled, void* dw1, void* db1, void* dw2, void* db2, int N, int C,
                          int Cs, cudaStream_t stream) {
  RT1_REQUIRE(Cs <= SE_MAX_CS && C <= 8192, "rt1_se_fc_bwd: unsupported C=%d Cs=%d", C, Cs);
  se_fc_bwd_kernel<<<N, SE_THREADS, (C + 2 * Cs) * sizeof(float), stream>>>(ds, s, z1, (const bf16*)w1, (const bf16*)w2, inv_hw, dz2, dz1,
                                                                       dpooled, C, Cs);
  int rc = rt1_check_launch("rt1_se_fc_bwd");
  if (rc) return rc;
  if (dw1 == nullptr) return 0;   // parameter gradients requested separately (rt1_se_wgrad), e.g. on another stream
  const long total = (long)C * Cs + C + Cs;
  se_wgrad_kernel<<<rt1_cdiv(total, 256), 256, 0, stream>>>(dz2, dz1, z1, pooled_sum, inv_hw, (bf16*)dw1, (bf16*)db1, (bf16*)dw2,
                                                           (bf16*)db2, N, C, Cs);
  return rt1_check_launch("rt1_se_wgrad");
}

// The parameter-gradient half of rt1_se_fc_bwd on its own (pass dw1 = NULL there): a leaf of the backward graph.
RT1_API int rt1_se_wgrad(const float* dz2, const float* dz1, const float* z1, const float* pooled_sum, float inv_hw, void* dw1,
                         void* db1, void* dw2, void* db2, int N, int C, int Cs, cudaStream_t stream) {
  const long total = (long)C * Cs + C + Cs;
  se_wgrad_kernel<<<rt1_cdiv(total, 256), 256, 0, stream>>>(dz2, dz1, z1, pooled_sum, inv_hw, (bf16*)dw1, (bf16*)db1, (bf16*)dw2,
                                                           (bf16*)db2, N, C, Cs);
  return rt1_check_launch("rt1_se_wgrad");
}
