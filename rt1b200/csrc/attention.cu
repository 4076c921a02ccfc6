// Fused masked attention over the RT-1 token window (replaces 2 bmm + masked_fill + softmax +
// dropout of K12) — GENERIC-SHAPE kernels on CUDA cores.  RT-1's own shapes (S <= 128, dk 64/128) are dispatched to the
// tcgen05 kernels in attention_tc.cu; these serve every other head size / window (and RT1B200_ATTN_TC=0 for A/B runs).
//
// The whole problem of one (batch, head) — S = 66 tokens, dk = 128 — fits in one CTA's shared memory, so the
// scores never touch HBM (the reference materialises [b, h, S, S] three times per layer).  Input is the fused
// projection output qkv [b, S, 3, h, dk] (bf16); output is [b, S, h*dk] ready for the out-projection GEMM.
//   forward : S = QK^T/sqrt(dk) → static RT-1 mask (0 → -1e9) → softmax (fp32) → dropout (stateless hash RNG,
//             regenerated in backward) → O = P V
//   backward: recomputes P from Q, K (nothing but qkv is saved), then dV, dP, dS, dQ, dK in the same CTA.
// Rows are padded to dk+8 bf16 in shared memory so 16-byte row reads by consecutive threads are conflict-free.
#include "common.cuh"

#define ATT_THREADS 1024   // one CTA per (batch, head) = 64 CTAs on 148 SMs at batch 8: 32 warps per CTA hide the shared-memory latency that 8 warps left exposed (ncu: 12 % occupancy, 47 % issue)

struct AttnDims {
  int B, S, H, DK;
  float scale;      // 1/sqrt(dk)
  float p_drop;
  uint32_t seed;
  const uint32_t* seed_epoch;   // device word added to `seed` (or null), see rt1_seed_epoch_ptr()
};
__device__ __forceinline__ uint32_t attn_seed(const AttnDims& d) { return d.seed + (d.seed_epoch != nullptr ? __ldg(d.seed_epoch) : 0u); }

__device__ __forceinline__ void load_tile(bf16* dst, const bf16* src, int S, int DK, int ld_src, int ld_dst) {
  const int vec_per_row = DK / 8;
  for (int i = threadIdx.x; i < S * vec_per_row; i += blockDim.x) {
    const int r = i / vec_per_row, c = i % vec_per_row;
    *reinterpret_cast<uint4*>(dst + r * ld_dst + c * 8) = *reinterpret_cast<const uint4*>(src + (size_t)r * ld_src + c * 8);
  }
}

__device__ __forceinline__ float dot_rows(const bf16* a, const bf16* b, int DK) {
  float acc = 0.f;
  for (int k = 0; k < DK; k += 8) {
    float fa[8], fb[8];
    unpack8(*reinterpret_cast<const bf16x8*>(a + k), fa);
    unpack8(*reinterpret_cast<const bf16x8*>(b + k), fb);
#pragma unroll
    for (int t = 0; t < 8; ++t) acc = fmaf(fa[t], fb[t], acc);
  }
  return acc;
}

// scores → masked softmax probabilities P (pre-dropout) in `prob` [S][S+1]
__device__ __forceinline__ void scores_softmax(const bf16* q, const bf16* k, float* prob, const uint8_t* mask, int S, int DK,
                                               int ld, float scale) {
  const int ldp = S + 1;
  for (int p = threadIdx.x; p < S * S; p += blockDim.x) {
    const int i = p / S, j = p % S;
    float s = dot_rows(q + i * ld, k + j * ld, DK) * scale;
    if (mask != nullptr && mask[i * S + j] == 0) s = -1e9f;
    prob[i * ldp + j] = s;
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  for (int i = warp; i < S; i += nw) {
    float mx = -INFINITY;
    for (int j = lane; j < S; j += 32) mx = fmaxf(mx, prob[i * ldp + j]);
    mx = warp_max(mx);
    float sum = 0.f;
    for (int j = lane; j < S; j += 32) {
      const float e = __expf(prob[i * ldp + j] - mx);
      prob[i * ldp + j] = e;
      sum += e;
    }
    sum = warp_sum(sum);
    const float inv = 1.f / sum;
    for (int j = lane; j < S; j += 32) prob[i * ldp + j] *= inv;
  }
  __syncthreads();
}

__global__ void __launch_bounds__(ATT_THREADS) attn_fwd_kernel(const bf16* __restrict__ qkv, const uint8_t* __restrict__ mask,
                                                               bf16* __restrict__ out, AttnDims d) {
  pdl_enter();
  extern __shared__ __align__(16) uint8_t smem[];
  const int S = d.S, DK = d.DK, ld = DK + 8, ldp = S + 1;
  bf16* q = reinterpret_cast<bf16*>(smem);
  bf16* k = q + S * ld;
  bf16* v = k + S * ld;
  float* prob = reinterpret_cast<float*>(v + S * ld);
  const int h = blockIdx.x, b = blockIdx.y;
  const int tok_stride = 3 * d.H * DK;
  const bf16* base = qkv + (size_t)b * S * tok_stride + (size_t)h * DK;
  load_tile(q, base, S, DK, tok_stride, ld);
  load_tile(k, base + (size_t)d.H * DK, S, DK, tok_stride, ld);
  load_tile(v, base + (size_t)2 * d.H * DK, S, DK, tok_stride, ld);
  __syncthreads();
  scores_softmax(q, k, prob, mask, S, DK, ld, d.scale);
  if (d.p_drop > 0.f) {
    const uint32_t thresh = rt1_drop_threshold(d.p_drop);
    const float ks = 1.f / (1.f - d.p_drop);
    const uint32_t head_base = (uint32_t)((b * d.H + h) * S * S);
    for (int p = threadIdx.x; p < S * S; p += blockDim.x) {
      const int i = p / S, j = p % S;
      const float pv = prob[i * ldp + j];
      prob[i * ldp + j] = rt1_hash_u32(head_base + (uint32_t)p, attn_seed(d)) >= thresh ? pv * ks : 0.f;
    }
    __syncthreads();
  }
  // O = P V ; task = (row i, 8-wide column chunk)
  const int chunks = DK / 8;
  bf16* obase = out + (size_t)b * S * d.H * DK + (size_t)h * DK;
  for (int t = threadIdx.x; t < S * chunks; t += blockDim.x) {
    const int i = t / chunks, c = t % chunks;
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int j = 0; j < S; ++j) {
      const float pv = prob[i * ldp + j];
      float fv[8];
      unpack8(*reinterpret_cast<const bf16x8*>(v + j * ld + c * 8), fv);
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] = fmaf(pv, fv[e], acc[e]);
    }
    *reinterpret_cast<bf16x8*>(obase + (size_t)i * d.H * DK + c * 8) = pack8(acc);
  }
}

__global__ void __launch_bounds__(ATT_THREADS) attn_bwd_kernel(const bf16* __restrict__ qkv, const uint8_t* __restrict__ mask,
                                                               const bf16* __restrict__ dout, bf16* __restrict__ dqkv,
                                                               AttnDims d) {
  pdl_enter();
  extern __shared__ __align__(16) uint8_t smem[];
  const int S = d.S, DK = d.DK, ld = DK + 8, ldp = S + 1;
  bf16* q = reinterpret_cast<bf16*>(smem);
  bf16* k = q + S * ld;
  bf16* v = k + S * ld;
  bf16* go = v + S * ld;
  float* prob = reinterpret_cast<float*>(go + S * ld);  // P (pre-dropout)
  float* ds = prob + S * ldp;                           // dP, then dS
  const int h = blockIdx.x, b = blockIdx.y;
  const int tok_stride = 3 * d.H * DK;
  const size_t off = (size_t)b * S * tok_stride + (size_t)h * DK;
  load_tile(q, qkv + off, S, DK, tok_stride, ld);
  load_tile(k, qkv + off + (size_t)d.H * DK, S, DK, tok_stride, ld);
  load_tile(v, qkv + off + (size_t)2 * d.H * DK, S, DK, tok_stride, ld);
  load_tile(go, dout + (size_t)b * S * d.H * DK + (size_t)h * DK, S, DK, d.H * DK, ld);
  __syncthreads();
  scores_softmax(q, k, prob, mask, S, DK, ld, d.scale);

  const uint32_t thresh = rt1_drop_threshold(d.p_drop);
  const float ks = d.p_drop > 0.f ? 1.f / (1.f - d.p_drop) : 1.f;
  const uint32_t head_base = (uint32_t)((b * d.H + h) * S * S);
  // dP[i][j] = keep * ks * <dO_i, V_j>
  for (int p = threadIdx.x; p < S * S; p += blockDim.x) {
    const int i = p / S, j = p % S;
    float g = dot_rows(go + i * ld, v + j * ld, DK);
    if (d.p_drop > 0.f) g = rt1_hash_u32(head_base + (uint32_t)p, attn_seed(d)) >= thresh ? g * ks : 0.f;
    ds[i * ldp + j] = g;
  }
  __syncthreads();
  const int chunks = DK / 8;
  bf16* dq_base = dqkv + off;
  bf16* dk_base = dq_base + (size_t)d.H * DK;
  bf16* dv_base = dq_base + (size_t)2 * d.H * DK;
  // dV[j][:] = sum_i Pd[i][j] dO[i][:]   (Pd = dropped-out probabilities)
  for (int t = threadIdx.x; t < S * chunks; t += blockDim.x) {
    const int j = t / chunks, c = t % chunks;
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int i = 0; i < S; ++i) {
      float pv = prob[i * ldp + j];
      if (d.p_drop > 0.f) pv = rt1_hash_u32(head_base + (uint32_t)(i * S + j), attn_seed(d)) >= thresh ? pv * ks : 0.f;
      float fo[8];
      unpack8(*reinterpret_cast<const bf16x8*>(go + i * ld + c * 8), fo);
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] = fmaf(pv, fo[e], acc[e]);
    }
    *reinterpret_cast<bf16x8*>(dv_base + (size_t)j * tok_stride + c * 8) = pack8(acc);
  }
  // dS = P * (dP - rowsum(dP * P)) * scale
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  for (int i = warp; i < S; i += nw) {
    float dotp = 0.f;
    for (int j = lane; j < S; j += 32) dotp += ds[i * ldp + j] * prob[i * ldp + j];
    dotp = warp_sum(dotp);
    for (int j = lane; j < S; j += 32) ds[i * ldp + j] = prob[i * ldp + j] * (ds[i * ldp + j] - dotp) * d.scale;
  }
  __syncthreads();
  // dQ[i][:] = sum_j dS[i][j] K[j][:] ; dK[j][:] = sum_i dS[i][j] Q[i][:]
  for (int t = threadIdx.x; t < 2 * S * chunks; t += blockDim.x) {
    const int which = t / (S * chunks), r = (t % (S * chunks)) / chunks, c = t % chunks;
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    const bf16* src = which == 0 ? k : q;
    for (int o = 0; o < S; ++o) {
      const float w = which == 0 ? ds[r * ldp + o] : ds[o * ldp + r];
      float f[8];
      unpack8(*reinterpret_cast<const bf16x8*>(src + o * ld + c * 8), f);
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] = fmaf(w, f[e], acc[e]);
    }
    bf16* dst = (which == 0 ? dq_base : dk_base) + (size_t)r * tok_stride + c * 8;
    *reinterpret_cast<bf16x8*>(dst) = pack8(acc);
  }
}

static size_t attn_smem(int S, int DK, bool bwd) {
  const size_t tile = (size_t)S * (DK + 8) * 2, pm = (size_t)S * (S + 1) * 4;
  return bwd ? 4 * tile + 2 * pm : 3 * tile + pm;
}

// tensor-core path (attention_tc.cu): S <= 128, dk in {64, 128}
bool rt1_attention_tc_supported(int S, int DK, const void* a, const void* b);
int rt1_attention_tc_fwd(const void* qkv, const unsigned char* mask, void* out, int B, int S, int H, int DK, float p_drop,
                         unsigned seed, cudaStream_t stream);
int rt1_attention_tc_bwd(const void* qkv, const unsigned char* mask, const void* dout, void* dqkv, int B, int S, int H, int DK,
                         float p_drop, unsigned seed, cudaStream_t stream);

RT1_API int rt1_attention_fwd(const void* qkv, const unsigned char* mask, void* out, int B, int S, int H, int DK, float p_drop,
                              unsigned seed, cudaStream_t stream) {
  if (rt1_attention_tc_supported(S, DK, qkv, out)) return rt1_attention_tc_fwd(qkv, mask, out, B, S, H, DK, p_drop, seed, stream);
  RT1_REQUIRE(DK % 8 == 0 && DK <= 256 && S <= 160, "rt1_attention_fwd: unsupported S=%d dk=%d", S, DK);
  const size_t smem = attn_smem(S, DK, false);
  RT1_REQUIRE(smem <= 227 * 1024, "rt1_attention_fwd: S=%d dk=%d needs %zu B of shared memory", S, DK, smem);
  if (int rc = rt1_ensure_dyn_smem((const void*)attn_fwd_kernel, smem)) return rc;
  AttnDims d{B, S, H, DK, 1.f / sqrtf((float)DK), p_drop, seed, p_drop > 0.f ? rt1_seed_epoch_ptr() : nullptr};
  rt1_launch(attn_fwd_kernel, dim3(H, B), ATT_THREADS, smem, stream, (const bf16*)qkv, mask, (bf16*)out, d);
  return rt1_check_launch("rt1_attention_fwd");
}

RT1_API int rt1_attention_bwd(const void* qkv, const unsigned char* mask, const void* dout, void* dqkv, int B, int S, int H,
                              int DK, float p_drop, unsigned seed, cudaStream_t stream) {
  if (rt1_attention_tc_supported(S, DK, qkv, dout) && ((uintptr_t)dqkv & 15) == 0)
    return rt1_attention_tc_bwd(qkv, mask, dout, dqkv, B, S, H, DK, p_drop, seed, stream);
  RT1_REQUIRE(DK % 8 == 0 && DK <= 256 && S <= 160, "rt1_attention_bwd: unsupported S=%d dk=%d", S, DK);
  const size_t smem = attn_smem(S, DK, true);
  RT1_REQUIRE(smem <= 227 * 1024, "rt1_attention_bwd: S=%d dk=%d needs %zu B of shared memory", S, DK, smem);
  if (int rc = rt1_ensure_dyn_smem((const void*)attn_bwd_kernel, smem)) return rc;
  AttnDims d{B, S, H, DK, 1.f / sqrtf((float)DK), p_drop, seed, p_drop > 0.f ? rt1_seed_epoch_ptr() : nullptr};
  rt1_launch(attn_bwd_kernel, dim3(H, B), ATT_THREADS, smem, stream, (const bf16*)qkv, mask, (const bf16*)dout, (bf16*)dqkv, d);
  return rt1_check_launch("rt1_attention_bwd");
}

// dropout mask re-application for the backward of a GEMM-epilogue dropout: out = keep ? in/(1-p) : 0
__global__ void dropout_apply_kernel(const bf16* __restrict__ in, bf16* __restrict__ out, long n, uint32_t thresh, float ks,
                                     uint32_t seed, const uint32_t* seed_epoch) {
  pdl_enter();
  if (seed_epoch != nullptr) seed += __ldg(seed_epoch);
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = rt1_hash_u32((uint32_t)i, seed) >= thresh ? __float2bfloat16_rn(__bfloat162float(in[i]) * ks) : __float2bfloat16_rn(0.f);
}
RT1_API int rt1_dropout_apply(const void* in, void* out, long long n, float p, unsigned seed, cudaStream_t stream) {
  rt1_launch(dropout_apply_kernel, rt1_cdiv(n, 256), 256, 0, stream, (const bf16*)in, (bf16*)out, n, rt1_drop_threshold(p), 1.f / (1.f - p), seed, rt1_seed_epoch_ptr());
  return rt1_check_launch("rt1_dropout_apply");
}
