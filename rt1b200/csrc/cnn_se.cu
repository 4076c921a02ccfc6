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
  pdl_enter();
  extern __shared__ __align__(16) float sm[];  // pooled [C] | hid [Cs]
  float* pooled = sm;
  float* hid = sm + C;
  const int n = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  for (int c = threadIdx.x; c < C; c += blockDim.x) pooled[c] = pooled_sum[(size_t)n * C + c] * inv_hw;
  __syncthreads();
  // a warp owns rows warp, warp + nw, ...: their loads are issued together (the row-after-row loop waited on one L2 round trip
  // per 256 channels and row — three rows deep at Cs = 96)
  constexpr int RPW = SE_MAX_CS / 32;
  {
    float acc[RPW];
#pragma unroll
    for (int j = 0; j < RPW; ++j) acc[j] = 0.f;
#pragma unroll 2
    for (int c = lane * 8; c < C; c += 256) {
      bf16x8 wv[RPW];
#pragma unroll
      for (int j = 0; j < RPW; ++j) {
        const int u = warp + j * nw;
        if (u < Cs) wv[j] = *reinterpret_cast<const bf16x8*>(w1 + (size_t)u * C + c);
      }
      const float4 p0 = *reinterpret_cast<const float4*>(pooled + c), p1 = *reinterpret_cast<const float4*>(pooled + c + 4);
      const float pv[8] = {p0.x, p0.y, p0.z, p0.w, p1.x, p1.y, p1.z, p1.w};
#pragma unroll
      for (int j = 0; j < RPW; ++j) {
        if (warp + j * nw < Cs) {
          float fw[8];
          unpack8(wv[j], fw);
#pragma unroll
          for (int e = 0; e < 8; ++e) acc[j] = fmaf(fw[e], pv[e], acc[j]);
        }
      }
    }
#pragma unroll
    for (int j = 0; j < RPW; ++j) {
      const int u = warp + j * nw;
      if (u < Cs) {                                   // warp-uniform
        const float a = warp_sum(acc[j]);
        if (lane == 0) {
          const float z = a + __bfloat162float(b1[u]);
          z1_out[(size_t)n * Cs + u] = z;
          hid[u] = silu_f(z);
        }
      }
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const bf162* wr = reinterpret_cast<const bf162*>(w2 + (size_t)c * Cs);   // Cs is even: rows are 4-byte aligned
    float a0 = __bfloat162float(b2[c]), a1 = 0.f, a2 = 0.f, a3 = 0.f;       // four chains: the adds do not serialise the loads
    int u = 0;
#pragma unroll 4
    for (; u + 8 <= Cs; u += 8) {
      const float2 w0 = __bfloat1622float2(wr[(u >> 1)]), w1v = __bfloat1622float2(wr[(u >> 1) + 1]);
      const float2 w2v = __bfloat1622float2(wr[(u >> 1) + 2]), w3 = __bfloat1622float2(wr[(u >> 1) + 3]);
      a0 = fmaf(w0.x, hid[u], a0);
      a0 = fmaf(w0.y, hid[u + 1], a0);
      a1 = fmaf(w1v.x, hid[u + 2], a1);
      a1 = fmaf(w1v.y, hid[u + 3], a1);
      a2 = fmaf(w2v.x, hid[u + 4], a2);
      a2 = fmaf(w2v.y, hid[u + 5], a2);
      a3 = fmaf(w3.x, hid[u + 6], a3);
      a3 = fmaf(w3.y, hid[u + 7], a3);
    }
    for (; u < Cs; u += 2) {
      const float2 w = __bfloat1622float2(wr[u >> 1]);
      a0 = fmaf(w.x, hid[u], a0);
      a1 = fmaf(w.y, hid[u + 1], a1);
    }
    const float acc = (a0 + a1) + (a2 + a3);
    s_out[(size_t)n * C + c] = sigmoid_f(acc);
  }
}

// ---------------------------------------------------------------------------------------------------- backward
// ds [N, C] is the gradient w.r.t. s.  With dz2 = ds s (1 - s), hid = SiLU(z1):
//   dhid[n,u]    = sum_c dz2[n,c] W2[c,u]      dz1 = dhid SiLU'(z1)      dpooled[n,c] = (1/HW) sum_u dz1[n,u] W1[u,c]
//   dW2[c,u] = sum_n dz2[n,c] hid[n,u]   db2[c] = sum_n dz2[n,c]   dW1[u,c] = sum_n dz1[n,u] pooled[n,c]   db1[u] = sum_n dz1[n,u]
// The first version ran one CTA per FRAME (each re-reading both weight matrices, 0.9 MB at C = 2304, from L2) plus a
// thread-per-weight gradient kernel: 30 + 26 us per block for 40 MFLOP.  Here a CTA owns a 32-CHANNEL SLICE for ALL frames,
// so every weight is read once per kernel, and the only cross-slice quantity — dhid — is summed with 16-byte fp32 reductions.
// Two kernels because dz1 needs the complete dhid:
//   A: dz2 slice → partial dhid (red.v4), dW2 rows and db2 of the slice
//   B: dz1 (every CTA, from the complete dhid) → dpooled slice, dW1 columns of the slice, db1 (CTA 0), and — the slice holding
//      s, dpooled and the A1..A4 parts of rt1_se_bn_bwd_reduce for all frames — the [2, C] sums of the BatchNorm backward
//      (previously rt1_bn_sums_from_parts, a launch of its own).
#define SE_SL 32
#define SE_BWD_THREADS 256

__device__ __forceinline__ void red_add_f32x4(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

__global__ void __launch_bounds__(SE_BWD_THREADS) se_bwd_a_kernel(const float* __restrict__ ds, const float* __restrict__ s,
                                                                  const float* __restrict__ z1, const bf16* __restrict__ w2,
                                                                  float* __restrict__ dhid, bf16* __restrict__ dw2, bf16* __restrict__ db2,
                                                                  int N, int C, int Cs) {
  pdl_enter();
  extern __shared__ __align__(16) float se_sm[];
  const int Cs4 = (Cs + 3) & ~3, U4 = Cs4 >> 2;
  float* hid = se_sm;                       // [N][Cs4]
  float* w2s = hid + N * Cs4;               // [32][Cs4]
  float* dz2s = w2s + SE_SL * Cs4;          // [N][33]  (odd row stride: rows of different frames fall into different banks)
  const int c0 = blockIdx.x * SE_SL, tid = threadIdx.x;
  for (int i = tid; i < N * SE_SL; i += SE_BWD_THREADS) {
    const int n = i / SE_SL, cl = i % SE_SL, c = c0 + cl;
    float g = 0.f;
    if (c < C) {
      const float sv = s[(size_t)n * C + c];
      g = ds[(size_t)n * C + c] * sv * (1.f - sv);
    }
    dz2s[n * (SE_SL + 1) + cl] = g;
  }
  for (int i = tid; i < N * Cs4; i += SE_BWD_THREADS) {
    const int n = i / Cs4, u = i - n * Cs4;
    hid[i] = u < Cs ? silu_f(z1[(size_t)n * Cs + u]) : 0.f;
  }
  for (int i = tid; i < SE_SL * Cs4; i += SE_BWD_THREADS) {
    const int cl = i / Cs4, u = i - cl * Cs4;
    w2s[i] = (u < Cs && c0 + cl < C) ? __bfloat162float(w2[(size_t)(c0 + cl) * Cs + u]) : 0.f;
  }
  __syncthreads();
  // partial dhid[n][4q .. 4q+3] over this slice: lanes walk q (consecutive 16-byte weight chunks), dz2 is a broadcast
  for (int it = tid; it < N * U4; it += SE_BWD_THREADS) {
    const int n = it / U4, q = it - n * U4;
    const float* dzr = dz2s + n * (SE_SL + 1);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 8
    for (int cl = 0; cl < SE_SL; ++cl) {
      const float g = dzr[cl];
      const float4 w = *reinterpret_cast<const float4*>(w2s + cl * Cs4 + 4 * q);
      acc.x = fmaf(g, w.x, acc.x);
      acc.y = fmaf(g, w.y, acc.y);
      acc.z = fmaf(g, w.z, acc.z);
      acc.w = fmaf(g, w.w, acc.w);
    }
    red_add_f32x4(dhid + (size_t)n * Cs4 + 4 * q, acc.x, acc.y, acc.z, acc.w);
  }
  // dW2[c][4q .. 4q+3]: lanes walk the slice's channels, hid is a 16-byte broadcast
  for (int it = tid; it < U4 * SE_SL; it += SE_BWD_THREADS) {
    const int cl = it % SE_SL, q = it / SE_SL;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
    for (int n = 0; n < N; ++n) {
      const float g = dz2s[n * (SE_SL + 1) + cl];
      const float4 h = *reinterpret_cast<const float4*>(hid + n * Cs4 + 4 * q);
      acc.x = fmaf(g, h.x, acc.x);
      acc.y = fmaf(g, h.y, acc.y);
      acc.z = fmaf(g, h.z, acc.z);
      acc.w = fmaf(g, h.w, acc.w);
    }
    const int c = c0 + cl;
    if (c < C) {
      const float v[4] = {acc.x, acc.y, acc.z, acc.w};
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (4 * q + k < Cs) dw2[(size_t)c * Cs + 4 * q + k] = __float2bfloat16_rn(v[k]);
    }
  }
  if (tid < SE_SL && c0 + tid < C) {
    float acc = 0.f;
    for (int n = 0; n < N; ++n) acc += dz2s[n * (SE_SL + 1) + tid];
    db2[c0 + tid] = __float2bfloat16_rn(acc);
  }
}

__global__ void __launch_bounds__(SE_BWD_THREADS) se_bwd_b_kernel(const float* __restrict__ dhid, const float* __restrict__ z1,
                                                                  const float* __restrict__ pooled_sum, float inv_hw,
                                                                  const bf16* __restrict__ w1, const float* __restrict__ parts,
                                                                  const float* __restrict__ s, float* __restrict__ dpooled,
                                                                  float* __restrict__ sums, bf16* __restrict__ dw1, bf16* __restrict__ db1,
                                                                  int N, int C, int Cs) {
  pdl_enter();
  extern __shared__ __align__(16) float se_sm[];
  const int Cs4 = (Cs + 3) & ~3, U4 = Cs4 >> 2, Np = (N + 3) & ~3, N4 = Np >> 2;
  float* dz1t = se_sm;                      // [Cs4][Np]  (transposed: 16-byte broadcasts along the frames)
  float* w1s = dz1t + Cs4 * Np;             // [Cs4][32]
  float* pl = w1s + Cs4 * SE_SL;            // [Np][32]  pooled means of the slice
  float* dps = pl + Np * SE_SL;             // [N][33]   dpooled of the slice
  float* red = dps + N * (SE_SL + 1);       // [2][256]  cross-lane reduction of the sums
  const int c0 = blockIdx.x * SE_SL, tid = threadIdx.x;
  for (int i = tid; i < Cs4 * Np; i += SE_BWD_THREADS) {
    const int n = i / Cs4, u = i - n * Cs4;                    // global reads contiguous along u
    float v = 0.f;
    if (n < N && u < Cs) v = __ldcg(dhid + (size_t)n * Cs4 + u) * silu_grad_f(z1[(size_t)n * Cs + u]);
    dz1t[u * Np + n] = v;
  }
  for (int i = tid; i < Cs4 * SE_SL; i += SE_BWD_THREADS) {
    const int u = i / SE_SL, cl = i % SE_SL;
    w1s[i] = (u < Cs && c0 + cl < C) ? __bfloat162float(w1[(size_t)u * C + c0 + cl]) : 0.f;
  }
  for (int i = tid; i < Np * SE_SL; i += SE_BWD_THREADS) {
    const int n = i / SE_SL, cl = i % SE_SL;
    pl[i] = (n < N && c0 + cl < C) ? pooled_sum[(size_t)n * C + c0 + cl] * inv_hw : 0.f;
  }
  __syncthreads();
  // dpooled[4g .. 4g+3][c]: lanes walk the channels of the slice
  for (int it = tid; it < N4 * SE_SL; it += SE_BWD_THREADS) {
    const int cl = it % SE_SL, g = it / SE_SL;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
    for (int u = 0; u < Cs; ++u) {
      const float w = w1s[u * SE_SL + cl];
      const float4 d = *reinterpret_cast<const float4*>(dz1t + u * Np + 4 * g);
      acc.x = fmaf(d.x, w, acc.x);
      acc.y = fmaf(d.y, w, acc.y);
      acc.z = fmaf(d.z, w, acc.z);
      acc.w = fmaf(d.w, w, acc.w);
    }
    const float v[4] = {acc.x * inv_hw, acc.y * inv_hw, acc.z * inv_hw, acc.w * inv_hw};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int n = 4 * g + k;
      if (n < N) {
        dps[n * (SE_SL + 1) + cl] = v[k];
        if (c0 + cl < C) dpooled[(size_t)n * C + c0 + cl] = v[k];
      }
    }
  }
  // dW1[4q .. 4q+3][c]
  for (int it = tid; it < U4 * SE_SL; it += SE_BWD_THREADS) {
    const int cl = it % SE_SL, q = it / SE_SL;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    const float* d0 = dz1t + (4 * q) * Np;
#pragma unroll 4
    for (int n = 0; n < N; ++n) {
      const float pv = pl[n * SE_SL + cl];
#pragma unroll
      for (int k = 0; k < 4; ++k) acc[k] = fmaf(d0[k * Np + n], pv, acc[k]);
    }
    if (c0 + cl < C) {
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (4 * q + k < Cs) dw1[(size_t)(4 * q + k) * C + c0 + cl] = __float2bfloat16_rn(acc[k]);
    }
  }
  if (blockIdx.x == 0 && tid < Cs) {
    float acc = 0.f;
    for (int n = 0; n < N; ++n) acc += dz1t[tid * Np + n];
    db1[tid] = __float2bfloat16_rn(acc);
  }
  if (parts == nullptr) return;
  __syncthreads();                                              // dps complete
  // sums[0,c] = sum_n s A1 + dpooled A2 ; sums[1,c] = sum_n s A3 + dpooled A4
  {
    const int cl = tid % SE_SL, grp = tid / SE_SL, c = c0 + cl;    // 8 frame lanes per channel
    const size_t plane = (size_t)N * C;
    float s1 = 0.f, s2 = 0.f;
    if (c < C)
#pragma unroll 3
      for (int n = grp; n < N; n += SE_BWD_THREADS / SE_SL) {
        const size_t i = (size_t)n * C + c;
        const float sv = s[i], dp = dps[n * (SE_SL + 1) + cl];
        s1 += sv * parts[plane + i] + dp * parts[2 * plane + i];
        s2 += sv * parts[3 * plane + i] + dp * parts[4 * plane + i];
      }
    red[tid] = s1;
    red[SE_BWD_THREADS + tid] = s2;
    __syncthreads();
    if (tid < SE_SL && c < C) {
      s1 = s2 = 0.f;
#pragma unroll
      for (int k = 0; k < SE_BWD_THREADS / SE_SL; ++k) {
        s1 += red[k * SE_SL + tid];
        s2 += red[SE_BWD_THREADS + k * SE_SL + tid];
      }
      sums[c] = s1;
      sums[C + c] = s2;
    }
  }
}

RT1_API int rt1_se_fc_fwd(const float* pooled_sum, float inv_hw, const void* w1, const void* b1, const void* w2, const void* b2,
                          float* z1, float* s, int N, int C, int Cs, cudaStream_t stream) {
  RT1_REQUIRE(Cs <= SE_MAX_CS && Cs % 2 == 0 && C % 8 == 0 && C <= 8192, "rt1_se_fc_fwd: unsupported C=%d Cs=%d", C, Cs);
  rt1_launch(se_fc_fwd_kernel, N, SE_THREADS, (C + Cs) * sizeof(float), stream, pooled_sum, inv_hw, (const bf16*)w1, (const bf16*)b1,
                                                                       (const bf16*)w2, (const bf16*)b2, z1, s, C, Cs);
  return rt1_check_launch("rt1_se_fc_fwd");
}

// Backward of the FC pair + its parameter gradients (+ the BatchNorm-backward sums of the tensor the SE scale multiplies).
//   ds, s fp32 [N,C]; z1 fp32 [N,Cs]; pooled_sum fp32 [N,C] (the forward's input); parts = out5 of rt1_se_bn_bwd_reduce or NULL
//   dhid: fp32 [N, ceil4(Cs)] scratch, ZEROED BY THE CALLER;  dpooled fp32 [N,C] (per-pixel gradient, already / HW);
//   sums fp32 [2,C] (written when parts != NULL: pass to rt1_bn_act_bwd with sums_ready = 1);  dw1 [Cs,C], db1 [Cs], dw2 [C,Cs],
//   db2 [C] bf16.
RT1_API int rt1_se_bwd(const float* ds, const float* s, const float* z1, const float* pooled_sum, float inv_hw, const void* w1,
                       const void* w2, const float* parts, float* dhid, float* dpooled, float* sums, void* dw1, void* db1, void* dw2,
                       void* db2, int N, int C, int Cs, cudaStream_t stream) {
  RT1_REQUIRE(N > 0 && C > 0 && Cs > 0 && Cs <= SE_BWD_THREADS, "rt1_se_bwd: unsupported N=%d C=%d Cs=%d", N, C, Cs);
  RT1_REQUIRE((parts == nullptr) == (sums == nullptr), "rt1_se_bwd: parts and sums go together");
  const int Cs4 = (Cs + 3) & ~3, Np = (N + 3) & ~3;
  const size_t smem_a = sizeof(float) * ((size_t)N * Cs4 + (size_t)SE_SL * Cs4 + (size_t)N * (SE_SL + 1));
  const size_t smem_b = sizeof(float) * ((size_t)Cs4 * Np + (size_t)Cs4 * SE_SL + (size_t)Np * SE_SL + (size_t)N * (SE_SL + 1) +
                                         2 * SE_BWD_THREADS);
  RT1_REQUIRE(smem_a <= 200 * 1024 && smem_b <= 200 * 1024, "rt1_se_bwd: N=%d Cs=%d need %zu / %zu bytes of shared memory", N, Cs,
              smem_a, smem_b);
  int rc;
  if ((rc = rt1_ensure_dyn_smem((const void*)se_bwd_a_kernel, 200 * 1024))) return rc;
  if ((rc = rt1_ensure_dyn_smem((const void*)se_bwd_b_kernel, 200 * 1024))) return rc;
  const int grid = (C + SE_SL - 1) / SE_SL;
  rt1_launch(se_bwd_a_kernel, grid, SE_BWD_THREADS, smem_a, stream, ds, s, z1, (const bf16*)w2, dhid, (bf16*)dw2, (bf16*)db2, N, C, Cs);
  if ((rc = rt1_check_launch("rt1_se_bwd(a)"))) return rc;
  rt1_launch(se_bwd_b_kernel, grid, SE_BWD_THREADS, smem_b, stream, dhid, z1, pooled_sum, inv_hw, (const bf16*)w1, parts, s, dpooled, sums,
                                                           (bf16*)dw1, (bf16*)db1, N, C, Cs);
  return rt1_check_launch("rt1_se_bwd(b)");
}
