// Shared pieces of the NHWC image-tokenizer kernels (cnn_pointwise.cu, cnn_dwconv.cu, cnn_se.cu).
//
// Layout: activations are bf16 [N, H*W, C] (NHWC), C a multiple of 8; statistics and per-channel / per-frame
// vectors are fp32.
//
// "Channel-stationary" thread mapping used by every elementwise / reduction kernel here: a CTA owns a range of
// rows (pixels) of ONE frame and a chunk of <= 256 channel-vectors (8 channels = 16 bytes each).  Thread t is
// bound to channel-vector cv = t % c8 for its whole life and walks rows r0 + k * rows_per_pass, so
//   * consecutive threads touch consecutive 16-byte chunks (rows are contiguous in memory → fully coalesced),
//   * per-channel parameters (BN scale/shift, mean/rstd, FiLM gamma/beta, SE scale) are loaded ONCE into
//     registers,
//   * per-channel reductions accumulate in registers and meet in shared memory only once at the end.
#pragma once
#include "common.cuh"

#define CNN_THREADS 256

struct ChanMap {
  int c8;             // channel vectors handled by this CTA (chunk size)
  int cv;             // this thread's channel vector inside the full tensor (global index, in units of 8 channels)
  int r0;             // first row offset of this thread inside the CTA's row range
  int rows_per_pass;  // row stride between the thread's consecutive rows
  bool active;
};

// C8 = C/8 of the full tensor; chunks along blockIdx.z
__device__ __forceinline__ ChanMap make_chan_map(int C8) {
  ChanMap m;
  const int nchunks = (C8 + 255) / 256;
  m.c8 = (C8 + nchunks - 1) / nchunks;
  const int chunk0 = blockIdx.z * m.c8;
  const int c8_here = min(m.c8, C8 - chunk0);
  m.rows_per_pass = CNN_THREADS / c8_here;
  const int t = threadIdx.x;
  m.active = t < m.rows_per_pass * c8_here;
  m.cv = chunk0 + t % c8_here;
  m.r0 = t / c8_here;
  m.c8 = c8_here;
  return m;
}
static inline int cnn_chan_chunks(int C) { return (C / 8 + 255) / 256; }

// Row range of a frame handled by this CTA (blockIdx.x of gridDim.x CTAs per frame; blockIdx.y = frame)
__device__ __forceinline__ void cta_rows(int HW, int& begin, int& end) {
  const int per = (HW + gridDim.x - 1) / gridDim.x;
  begin = blockIdx.x * per;
  end = min(HW, begin + per);
}
static inline int cnn_ctas_per_frame(int N, int HW, int C) {
  // aim for >= ~4 CTAs per SM overall, but keep at least 32 rows per CTA
  int want = (148 * 8 + N - 1) / N;
  int maxc = (HW + 31) / 32;
  if (want > maxc) want = maxc;
  if (want < 1) want = 1;
  return want;
}

// ---- packed fp32 pairs (fma.rn.f32x2 and friends): two channels per instruction -----------------------------------------
// The elementwise BatchNorm / SiLU passes are issue-bound on scalar FFMA / FMUL / FADD (ncu: 55 % of the executed instructions
// of se_bn_bwd_reduce); Blackwell executes the x2 forms at the same rate per instruction.
typedef unsigned long long p2;
__device__ __forceinline__ p2 p2_pack(float lo, float hi) {
  p2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void p2_unpack(p2 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ p2 p2_fma(p2 a, p2 b, p2 c) {
  p2 d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ p2 p2_mul(p2 a, p2 b) {
  p2 d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ p2 p2_add(p2 a, p2 b) {
  p2 d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
// sigmoid of a pair given u: tanh(u/2)/2 + 1/2 (two MUFU.TANH, two packed ops)
__device__ __forceinline__ p2 p2_sigmoid(p2 u) {
  const p2 half = p2_pack(0.5f, 0.5f);
  float h0, h1;
  p2_unpack(p2_mul(u, half), h0, h1);
  return p2_fma(p2_pack(tanh_approx_f(h0), tanh_approx_f(h1)), half, half);
}
// eight consecutive floats / one 16-byte bf16 vector as four pairs
__device__ __forceinline__ void load8p(const float* p, p2* f) {
  const float4 a = reinterpret_cast<const float4*>(p)[0], b = reinterpret_cast<const float4*>(p)[1];
  f[0] = p2_pack(a.x, a.y); f[1] = p2_pack(a.z, a.w); f[2] = p2_pack(b.x, b.y); f[3] = p2_pack(b.z, b.w);
}
__device__ __forceinline__ void unpack8p(const bf16x8& v, p2* f) {
  const uint32_t* w = reinterpret_cast<const uint32_t*>(&v);
#pragma unroll
  for (int j = 0; j < 4; ++j) f[j] = p2_pack(__uint_as_float(w[j] << 16), __uint_as_float(w[j] & 0xffff0000u));
}
__device__ __forceinline__ void p2_to8(const p2* v, float* f) {
#pragma unroll
  for (int j = 0; j < 4; ++j) p2_unpack(v[j], f[2 * j], f[2 * j + 1]);
}

__device__ __forceinline__ void load8f(const float* p, float* f) {
  const float4 a = reinterpret_cast<const float4*>(p)[0], b = reinterpret_cast<const float4*>(p)[1];
  f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}
__device__ __forceinline__ void load8bf(const bf16* p, float* f) { unpack8(*reinterpret_cast<const bf16x8*>(p), f); }
__device__ __forceinline__ void store8bf(bf16* p, const float* f) { *reinterpret_cast<bf16x8*>(p) = pack8(f); }

// Reduce per-thread 8-channel accumulators over the threads of a CTA that share a channel vector, then add the
// CTA totals to global memory with one atomic per channel.  `sm` needs CNN_THREADS * 8 floats.
__device__ __forceinline__ void cta_channel_reduce_atomic(const ChanMap& m, const float* acc, float* sm, float* gdst) {
  __syncthreads();
#pragma unroll
  for (int k = 0; k < 8; ++k) sm[threadIdx.x * 8 + k] = m.active ? acc[k] : 0.f;
  __syncthreads();
  // thread t < c8*8 sums channel (t / 8 vector, t % 8 lane) over the rows_per_pass threads that own it
  for (int t = threadIdx.x; t < m.c8 * 8; t += blockDim.x) {
    const int v = t >> 3, k = t & 7;
    float s = 0.f;
    for (int r = 0; r < m.rows_per_pass; ++r) s += sm[(r * m.c8 + v) * 8 + k];
    const int chunk0 = m.cv - (threadIdx.x % m.c8);  // first channel vector of this CTA's chunk (same for all threads)
    atomicAdd(gdst + (size_t)(chunk0 + v) * 8 + k, s);
  }
}
