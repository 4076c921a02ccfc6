// Frame pre-processing in one pass (replaces /255 + F.pad + advanced-index gather of K1).
//
// in : [N, 3, H, W] uint8 (0..255) or float32 (0..1), NCHW as the data loader delivers it
// out: [N, H, W, 8] bf16 NHWC, channels 3..7 zero (16-byte pixels: TMA- and vector-load friendly)
// out[n, y, x, c] = in[n, c, y + dh - pad_h, x + dw - pad_w] (zero outside) — the reference's
// zero-pad-by-7%-then-shifted-window crop with ONE offset for the whole batch (preprocessors.py:37-56).
#include "common.cuh"

template <typename T>
__global__ void __launch_bounds__(256) preprocess_kernel(const T* __restrict__ in, bf16* __restrict__ out, int N, int H, int W,
                                                         int off_y, int off_x, const int* __restrict__ shift_dev, float scale) {
  pdl_enter();
  if (shift_dev != nullptr) {   // offset sampled on the device (graph-captured steps): (dh, dw) added to (-pad_h, -pad_w)
    off_y += __ldg(shift_dev);
    off_x += __ldg(shift_dev + 1);
  }
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

static int preprocess_launch(const void* in, int in_is_uint8, void* out, int N, int H, int W, int off_y, int off_x,
                             const int* shift_dev, cudaStream_t stream) {
  const long total = (long)N * H * W;
  const int grid = rt1_cdiv(total, 256);
  if (in_is_uint8)
    rt1_launch(preprocess_kernel<uint8_t>, grid, 256, 0, stream, (const uint8_t*)in, (bf16*)out, N, H, W, off_y, off_x, shift_dev, 1.f / 255.f);
  else
    rt1_launch(preprocess_kernel<float>, grid, 256, 0, stream, (const float*)in, (bf16*)out, N, H, W, off_y, off_x, shift_dev, 1.f);
  return rt1_check_launch("rt1_preprocess_nhwc");
}

RT1_API int rt1_preprocess_nhwc(const void* in, int in_is_uint8, void* out, int N, int H, int W, int shift_h, int shift_w,
                                int pad_h, int pad_w, cudaStream_t stream) {
  return preprocess_launch(in, in_is_uint8, out, N, H, W, shift_h - pad_h, shift_w - pad_w, nullptr, stream);
}

// Same, with the (dh, dw) offset read from device memory (int32[2]) when the kernel runs: a CUDA graph bakes scalar
// arguments, so a captured training step samples the offset with a device-side RNG and passes its address.
RT1_API int rt1_preprocess_nhwc_dev(const void* in, int in_is_uint8, void* out, int N, int H, int W, const int* shift_dev,
                                    int pad_h, int pad_w, cudaStream_t stream) {
  RT1_REQUIRE(shift_dev != nullptr, "rt1_preprocess_nhwc_dev: shift_dev is null");
  return preprocess_launch(in, in_is_uint8, out, N, H, W, -pad_h, -pad_w, shift_dev, stream);
}
