// 256-bin action-token cross-entropy (replaces gather + ATen CE + argmax of K13).
//
// logits [R, A, V] bf16 (already gathered at the action positions), labels [R, A] int64.
// Forward: one warp per (r, a) row: max / log-sum-exp / argmax in fp32; one more pass folds the A rows of a
// sample into loss[r] = mean_a(CE) / num_items (the reference's batch-size dependent scaling,
// transformer_network.py:314-322).  Backward: dlogits = (softmax - onehot) * dloss[r] / (A * num_items).
#include "common.cuh"

template <int VPL>
__global__ void __launch_bounds__(256) ce_fwd_kernel(const bf16* __restrict__ logits, const long long* __restrict__ labels,
                                                     float* __restrict__ row_ce, float* __restrict__ row_lse,
                                                     long long* __restrict__ pred, int nrows, int V) {
  pdl_enter();
  const int lane = threadIdx.x & 31, row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= nrows) return;
  const bf16x8* lr = reinterpret_cast<const bf16x8*>(logits + (size_t)row * V);
  float v[VPL * 8];
  float mx = -INFINITY;
  int arg = 0;
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    unpack8(lr[lane + 32 * i], v + 8 * i);
#pragma unroll
    for (int k = 0; k < 8; ++k)
      if (v[8 * i + k] > mx) {
        mx = v[8 * i + k];
        arg = (lane + 32 * i) * 8 + k;
      }
  }
  // warp arg-max, ties → lowest index (torch.argmax semantics)
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    float om = __shfl_xor_sync(0xffffffffu, mx, o);
    int oa = __shfl_xor_sync(0xffffffffu, arg, o);
    if (om > mx || (om == mx && oa < arg)) {
      mx = om;
      arg = oa;
    }
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < VPL * 8; ++i) s += __expf(v[i] - mx);
  s = warp_sum(s);
  const float lse = mx + __logf(s);
  const long long lab = labels[row];
  // the label's logit: owned by exactly one lane
  float ll = 0.f;
#pragma unroll
  for (int i = 0; i < VPL; ++i)
#pragma unroll
    for (int k = 0; k < 8; ++k)
      if ((lane + 32 * i) * 8 + k == (int)lab) ll = v[8 * i + k];
  ll = warp_sum(ll);
  if (lane == 0) {
    row_ce[row] = lse - ll;
    row_lse[row] = lse;
    pred[row] = arg;
  }
}

__global__ void ce_fold_kernel(const float* __restrict__ row_ce, float* __restrict__ loss, int R, int A, float inv) {
  pdl_enter();
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= R) return;
  float s = 0.f;
  for (int a = 0; a < A; ++a) s += row_ce[r * A + a];
  loss[r] = s * inv;
}

template <int VPL>
__global__ void __launch_bounds__(256) ce_bwd_kernel(const bf16* __restrict__ logits, const long long* __restrict__ labels,
                                                     const float* __restrict__ row_lse, const float* __restrict__ dloss,
                                                     bf16* __restrict__ dlogits, int nrows, int A, int V, float inv) {
  pdl_enter();
  const int lane = threadIdx.x & 31, row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= nrows) return;
  const float scale = dloss[row / A] * inv, lse = row_lse[row];
  const int lab = (int)labels[row];
  const bf16x8* lr = reinterpret_cast<const bf16x8*>(logits + (size_t)row * V);
  bf16x8* dr = reinterpret_cast<bf16x8*>(dlogits + (size_t)row * V);
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    float v[8], o[8];
    unpack8(lr[lane + 32 * i], v);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int col = (lane + 32 * i) * 8 + k;
      o[k] = (__expf(v[k] - lse) - (col == lab ? 1.f : 0.f)) * scale;
    }
    dr[lane + 32 * i] = pack8(o);
  }
}

// row_ce / row_lse: fp32 [R*A] scratch that the backward reuses (row_lse).
RT1_API int rt1_action_ce_fwd(const void* logits, const long long* labels, float* loss, float* row_ce, float* row_lse,
                              long long* pred, int R, int A, int V, float num_items, cudaStream_t stream) {
  RT1_REQUIRE(V % 256 == 0 && V <= 1024, "rt1_action_ce_fwd: V=%d must be a multiple of 256 and <= 1024", V);
  const int nrows = R * A, wpb = 8, grid = rt1_cdiv(nrows, wpb);
  switch (V / 256) {
    case 1: rt1_launch(ce_fwd_kernel<1>, grid, wpb * 32, 0, stream, (const bf16*)logits, labels, row_ce, row_lse, pred, nrows, V); break;
    case 2: rt1_launch(ce_fwd_kernel<2>, grid, wpb * 32, 0, stream, (const bf16*)logits, labels, row_ce, row_lse, pred, nrows, V); break;
    case 4: rt1_launch(ce_fwd_kernel<4>, grid, wpb * 32, 0, stream, (const bf16*)logits, labels, row_ce, row_lse, pred, nrows, V); break;
    default: RT1_REQUIRE(false, "rt1_action_ce_fwd: unsupported V=%d", V);
  }
  int rc = rt1_check_launch("rt1_action_ce_fwd");
  if (rc) return rc;
  rt1_launch(ce_fold_kernel, rt1_cdiv(R, 128), 128, 0, stream, row_ce, loss, R, A, 1.f / ((float)A * num_items));
  return rt1_check_launch("rt1_action_ce_fold");
}

RT1_API int rt1_action_ce_bwd(const void* logits, const long long* labels, const float* row_lse, const float* dloss,
                              void* dlogits, int R, int A, int V, float num_items, cudaStream_t stream) {
  RT1_REQUIRE(V % 256 == 0 && V <= 1024, "rt1_action_ce_bwd: V=%d must be a multiple of 256 and <= 1024", V);
  const int nrows = R * A, wpb = 8, grid = rt1_cdiv(nrows, wpb);
  const float inv = 1.f / ((float)A * num_items);
  switch (V / 256) {
    case 1: rt1_launch(ce_bwd_kernel<1>, grid, wpb * 32, 0, stream, (const bf16*)logits, labels, row_lse, dloss, (bf16*)dlogits, nrows, A, V, inv); break;
    case 2: rt1_launch(ce_bwd_kernel<2>, grid, wpb * 32, 0, stream, (const bf16*)logits, labels, row_lse, dloss, (bf16*)dlogits, nrows, A, V, inv); break;
    case 4: rt1_launch(ce_bwd_kernel<4>, grid, wpb * 32, 0, stream, (const bf16*)logits, labels, row_lse, dloss, (bf16*)dlogits, nrows, A, V, inv); break;
    default: RT1_REQUIRE(false, "rt1_action_ce_bwd: unsupported V=%d", V);
  }
  return rt1_check_launch("rt1_action_ce_bwd");
}
