// AdamW update rule shared by the local (adamw.cu) and the NVLink-fused (p2p.cu) optimizer kernels.
#pragma once
#include "common.cuh"

struct AdamHyper {
  float lr, beta1, beta2, eps, wd, inv_bc1, inv_sqrt_bc2, grad_scale;
};

__device__ __forceinline__ void adam_update(float& p, float g, float& m, float& v, const AdamHyper& h) {
  g *= h.grad_scale;
  p *= (1.f - h.lr * h.wd);
  m = h.beta1 * m + (1.f - h.beta1) * g;
  v = h.beta2 * v + (1.f - h.beta2) * g * g;
  float denom = sqrtf(v) * h.inv_sqrt_bc2 + h.eps;
  p -= h.lr * h.inv_bc1 * (m / denom);
}

