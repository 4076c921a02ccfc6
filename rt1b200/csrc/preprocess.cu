// Frame pre-processing in one pass (replaces /255 + F.pad + advanced-index gather of K1).
//
// in : [N, 3, H, W] uint8 (0..255) or float32 (0..1), NCHW as the data loader delivers it
// out: [N, H, W, 8] bf16 NHWC, channels 3..7 zero (16-byte pixels: TMA- and vector-load friendly)
// out[n, y, x, c] = in[n, c, y + dh - pad_h, x + dw - pad_w] (zero outside) — the reference's
// zero-pad-by-7%-then-shifted-window crop with ONE offset for the whole batch (preprocessors.py:37-56).
#include "common.cuh"

template <typename T>
__global__ void __launch_bounds__(256) preprocess_kernel(const T* __restrict__ in, bf16* __restrict__ out, int N, int H, int W,
                                                         int off_y, int off_x, float scale) {
  const long total = (long)N * H * W;
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int x = (int)(i % W);
  const int y = (int)((i / W) % H);
  const long n = i / ((long)W * H);
  const int sy = y + off_y, sx = x + off_x;
  float c[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (sy >= 0 && sy < H && sx >= 0 && sx < W) {
    const T* p = in + (n * 3 * H + sy) * (long)W + sx;
    const long plane = (long)H * W;
    c[0] = (float)p[0] * scale;
    c[1] = (float)p[plane] * scale;
    c[2] = (float)p[2 * plane] * scale;
  }
  reinterpret_cast<bf16x8*>(out)[i] = pack8(c);
}

RT1_API int rt1_preprocess_nhwc(const void* in, int in_is_uint8, void* out, int N, int H, int W, int shift_h, int shift_w,
                                int pad_h, int pad_w, cudaStream_t stream) {
  const long total = (long)N * H * W;
  const int grid = rt1_cdiv(total, 256);
  if (in_is_uint8)
    preprocess_kernel<uint8_t><<<grid, 256, 0, stream>>>((const uint8_t*)in, (bf16*)out, N, H, W, shift_h - pad_h, shift_w - pad_w, 1.f / 255.f);
  else
    preprocess_kernel<float><<<grid, 256, 0, stream>>>((const float*)in, (bf16*)out, N, H, W, shift_h - pad_h, shift_w - pad_w, 1.f);
  return rt1_check_launch("rt1_preprocess_nhwc");
}
