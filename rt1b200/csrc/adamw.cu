// Fused flat-buffer AdamW (world size 1 path of the reduce+update step; SURVEY §8 A1, replaces K14).
//
// One pass over a contiguous bucket: g (bf16 or fp32) is scaled, the fp32 master weight and both moments are
// updated, the bf16 compute copy of the parameter is rewritten and the gradient slot is cleared for the next
// step.  30 B/element of HBM traffic, no per-tensor launches (the reference runs torch's foreach Adam over
// 572 tensors plus a separate zero_grad).  Decoupled weight decay; with wd == 0 this is exactly torch.optim.Adam.
#include "common.cuh"

#include "adam.cuh"

template <bool GRAD_BF16>
__global__ void __launch_bounds__(256) adamw_flat_kernel(bf16* __restrict__ p_lp, float* __restrict__ p, void* __restrict__ g_,
                                                         float* __restrict__ m, float* __restrict__ v, long n, AdamHyper h,
                                                         int zero_grad) {
  pdl_enter();
  const long n8 = n >> 3;
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const long stride = (long)gridDim.x * blockDim.x;
  for (; i < n8; i += stride) {
    float g[8];
    if (GRAD_BF16) {
      bf16x8 gv = reinterpret_cast<const bf16x8*>(g_)[i];
      unpack8(gv, g);
    } else {
      float4 a = reinterpret_cast<const float4*>(g_)[2 * i], b = reinterpret_cast<const float4*>(g_)[2 * i + 1];
      g[0] = a.x; g[1] = a.y; g[2] = a.z; g[3] = a.w; g[4] = b.x; g[5] = b.y; g[6] = b.z; g[7] = b.w;
    }
    float4 p0 = reinterpret_cast<float4*>(p)[2 * i], p1 = reinterpret_cast<float4*>(p)[2 * i + 1];
    float4 m0 = reinterpret_cast<float4*>(m)[2 * i], m1 = reinterpret_cast<float4*>(m)[2 * i + 1];
    float4 v0 = reinterpret_cast<float4*>(v)[2 * i], v1 = reinterpret_cast<float4*>(v)[2 * i + 1];
    float pp[8] = {p0.x, p0.y, p0.z, p0.w, p1.x, p1.y, p1.z, p1.w};
    float mm[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
    float vv[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
    for (int k = 0; k < 8; ++k) adam_update(pp[k], g[k], mm[k], vv[k], h);
    reinterpret_cast<float4*>(p)[2 * i] = make_float4(pp[0], pp[1], pp[2], pp[3]);
    reinterpret_cast<float4*>(p)[2 * i + 1] = make_float4(pp[4], pp[5], pp[6], pp[7]);
    reinterpret_cast<float4*>(m)[2 * i] = make_float4(mm[0], mm[1], mm[2], mm[3]);
    reinterpret_cast<float4*>(m)[2 * i + 1] = make_float4(mm[4], mm[5], mm[6], mm[7]);
    reinterpret_cast<float4*>(v)[2 * i] = make_float4(vv[0], vv[1], vv[2], vv[3]);
    reinterpret_cast<float4*>(v)[2 * i + 1] = make_float4(vv[4], vv[5], vv[6], vv[7]);
    if (p_lp) reinterpret_cast<bf16x8*>(p_lp)[i] = pack8(pp);
    if (zero_grad) {
      if (GRAD_BF16) {
        reinterpret_cast<uint4*>(g_)[i] = make_uint4(0, 0, 0, 0);
      } else {
        reinterpret_cast<float4*>(g_)[2 * i] = make_float4(0, 0, 0, 0);
        reinterpret_cast<float4*>(g_)[2 * i + 1] = make_float4(0, 0, 0, 0);
      }
    }
  }
  // scalar tail (buckets are 64-element aligned, so normally empty)
  for (long j = (n8 << 3) + (long)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += stride) {
    float g = GRAD_BF16 ? __bfloat162float(reinterpret_cast<bf16*>(g_)[j]) : reinterpret_cast<float*>(g_)[j];
    float pp = p[j], mm = m[j], vv = v[j];
    adam_update(pp, g, mm, vv, h);
    p[j] = pp; m[j] = mm; v[j] = vv;
    if (p_lp) p_lp[j] = __float2bfloat16_rn(pp);
    if (zero_grad) {
      if (GRAD_BF16) reinterpret_cast<bf16*>(g_)[j] = __float2bfloat16_rn(0.f);
      else reinterpret_cast<float*>(g_)[j] = 0.f;
    }
  }
}

RT1_API int rt1_fused_adamw(void* param_bf16, float* master, void* grad, int grad_is_bf16, float* exp_avg,
                            float* exp_avg_sq, long long n, float lr, float beta1, float beta2, float eps,
                            float weight_decay, float bias_corr1, float bias_corr2, float grad_scale, int zero_grad,
                            cudaStream_t stream) {
  if (n <= 0) return 0;
  RT1_REQUIRE(((uintptr_t)master % 16 == 0) && ((uintptr_t)grad % 16 == 0) && ((uintptr_t)exp_avg % 16 == 0) &&
                  ((uintptr_t)exp_avg_sq % 16 == 0) && ((uintptr_t)param_bf16 % 16 == 0),
              "rt1_fused_adamw: buffers must be 16-byte aligned");
  AdamHyper h{lr, beta1, beta2, eps, weight_decay, 1.f / bias_corr1, 1.f / sqrtf(bias_corr2), grad_scale};
  long n8 = n >> 3;
  int blocks = (int)((n8 + 255) / 256);
  if (blocks > 148 * 16) blocks = 148 * 16;
  if (blocks < 1) blocks = 1;
  if (grad_is_bf16)
    rt1_launch(adamw_flat_kernel<true>, blocks, 256, 0, stream, (bf16*)param_bf16, master, grad, exp_avg, exp_avg_sq, n, h, zero_grad);
  else
    rt1_launch(adamw_flat_kernel<false>, blocks, 256, 0, stream, (bf16*)param_bf16, master, grad, exp_avg, exp_avg_sq, n, h, zero_grad);
  return rt1_check_launch("rt1_fused_adamw");
}
