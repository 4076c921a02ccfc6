// Action (de)tokenisation (replaces the clamp / affine / cast chain of reference
// tokenizers/action_tokenizer.py:105-159).  One token per action dimension, ordered as the action space's keys:
//   Discrete(n):  token = value;                       detokenise: token > n ? 0 : token   ('>' as upstream, not '>=')
//   Box(low, high): token = int32(trunc((clamp(a) - low) / (high - low) * (vocab - 1)));  detokenise: the inverse affine map
// The caller packs the raw action values of a step as one fp32 row [A] (discrete values as floats); low/high/discrete are
// per-token tables (for a discrete token `high` holds n).
#include "common.cuh"

__global__ void action_tokenize_kernel(const float* __restrict__ vals, const float* __restrict__ low, const float* __restrict__ high,
                                       const uint8_t* __restrict__ discrete, long long* __restrict__ tokens, long total, int A,
                                       float vmax) {
  pdl_enter();
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int k = (int)(i % A);
  const float a = vals[i];
  long long t;
  if (discrete[k]) {
    t = (long long)a;
  } else {
    const float lo = low[k], hi = high[k];
    const float c = fminf(fmaxf(a, lo), hi);
    t = (long long)(int)((c - lo) / (hi - lo) * vmax);     // truncation, like the reference's int32 cast
  }
  tokens[i] = t;
}

__global__ void action_detokenize_kernel(const long long* __restrict__ tokens, const float* __restrict__ low,
                                         const float* __restrict__ high, const uint8_t* __restrict__ discrete,
                                         float* __restrict__ vals, long total, int A, float vmax) {
  pdl_enter();
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int k = (int)(i % A);
  const long long t = tokens[i];
  float v;
  if (discrete[k]) v = (float)(t > (long long)high[k] ? 0 : t);
  else v = (float)t / vmax * (high[k] - low[k]) + low[k];
  vals[i] = v;
}

// vals fp32 [R, A] → tokens int64 [R, A]
RT1_API int rt1_action_tokenize(const float* vals, const float* low, const float* high, const unsigned char* discrete,
                                long long* tokens, long long R, int A, int vocab, cudaStream_t stream) {
  const long total = (long)R * A;
  if (total <= 0) return 0;
  rt1_launch(action_tokenize_kernel, rt1_cdiv(total, 256), 256, 0, stream, vals, low, high, discrete, tokens, total, A, (float)(vocab - 1));
  return rt1_check_launch("rt1_action_tokenize");
}

// tokens int64 [R, A] → vals fp32 [R, A]
RT1_API int rt1_action_detokenize(const long long* tokens, const float* low, const float* high, const unsigned char* discrete,
                                  float* vals, long long R, int A, int vocab, cudaStream_t stream) {
  const long total = (long)R * A;
  if (total <= 0) return 0;
  rt1_launch(action_detokenize_kernel, rt1_cdiv(total, 256), 256, 0, stream, tokens, low, high, discrete, vals, total, A, (float)(vocab - 1));
  return rt1_check_launch("rt1_action_detokenize");
}
