// "Last CTA finalises": BatchNorm statistics → scale / shift inside the kernel that produced the statistics.
//
// Every producer of a BatchNorm input (the persistent GEMM, the depthwise tile kernel) adds its per-channel sum / sum of
// squares to global memory with atomics.  Turning those into a = gamma*rstd, b = beta - mean*a (+ running statistics) took
// a separate one-CTA launch per BatchNorm layer — 78 launches per training step that each wait for the producer to drain
// and delay its consumer.  With a tail, each CTA takes a ticket after its atomics; the CTA that draws the last ticket
// knows every contribution has landed and does the finalisation itself.  No CTA waits for another one.
#pragma once
#include "common.cuh"
#include "../../include/rt1b200.h"

// All threads of the CTA call this after the CTA's atomics to sum / sumsq were issued.  n_ctas = number of CTAs that call;
// s_last = one word of the caller's shared memory (the persistent GEMM has no static shared memory to spare: it asks for the
// full 227 KB as dynamic).
__device__ __forceinline__ void bn_tail_run(const Rt1BnTail& t, const float* sum, const float* sumsq, int C, unsigned n_ctas,
                                            unsigned* s_last) {
  __threadfence();                                   // this thread's atomics are ordered before the ticket
  __syncthreads();
  if (threadIdx.x == 0) *s_last = (atomicAdd(t.ticket, 1u) == n_ctas - 1u) ? 1u : 0u;
  __syncthreads();
  if (!*s_last) return;
  __threadfence();
  const bf16* __restrict__ gamma = reinterpret_cast<const bf16*>(t.gamma);
  const bf16* __restrict__ beta = reinterpret_cast<const bf16*>(t.beta);
  // one CTA does this alone while the rest of the GPU idles: every load of up to four channels per thread is issued before
  // the first is used (a channel-at-a-time loop paid two L2 round trips per channel, 7 us for 2304 channels)
  constexpr int U = 4;
  for (int c0 = threadIdx.x; c0 < C; c0 += U * blockDim.x) {
    float s1[U], s2[U], rm[U], rv[U], g[U], bt[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int c = c0 + u * blockDim.x;
      if (c < C) {
        s1[u] = __ldcg(sum + c);                      // L2 reads: the sums were built there by the atomics
        s2[u] = __ldcg(sumsq + c);
        g[u] = __bfloat162float(gamma[c]);
        bt[u] = __bfloat162float(beta[c]);
        if (t.running_mean != nullptr) {
          rm[u] = t.running_mean[c];
          rv[u] = t.running_var[c];
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int c = c0 + u * blockDim.x;
      if (c >= C) continue;
      const float mean = s1[u] / t.count;
      const float var = fmaxf(s2[u] / t.count - mean * mean, 0.f);
      if (t.running_mean != nullptr) {
        t.running_mean[c] = (1.f - t.momentum) * rm[u] + t.momentum * mean;
        const float unbiased = t.count > 1.f ? var * t.count / (t.count - 1.f) : var;
        t.running_var[c] = (1.f - t.momentum) * rv[u] + t.momentum * unbiased;
      }
      const float rstd = rsqrtf(var + t.eps);
      t.a[c] = g[u] * rstd;
      t.b[c] = bt[u] - mean * g[u] * rstd;
      t.mean[c] = mean;
      t.rstd[c] = rstd;
    }
  }
}
