// Batched z-buffered triangle rasteriser for the synthetic Language-Table board, plus the fused
// centre-crop + bilinear resize that turns rendered frames into policy inputs.
//
// What it replaces: PyBullet's TinyRenderer call per environment step (reference language_table.py:579-597, a CPU
// software rasteriser inside the pybullet wheel) and the TF crop/resize of the eval wrapper (eval/wrappers.py:88-96).
// No graphics API is involved: triangles arrive already projected to screen space (x, y in pixels, w = 1/depth),
// each CTA owns a 16x16 pixel tile of one environment, stages that environment's triangles through shared memory
// and keeps the nearest hit per pixel.  Output is a headless device buffer uint8 [E, H, W, 3].
#include "common.cuh"

#define RAST_TILE 16
#define RAST_CHUNK 128

struct RastTri {
  float x0, y0, w0, x1, y1, w1, x2, y2, w2;  // screen-space vertices, w = 1 / camera depth (> 0 in front of the camera)
  float r, g, b;
};

__global__ void __launch_bounds__(RAST_TILE * RAST_TILE) raster_kernel(const float* __restrict__ tris, const int* __restrict__ ntris,
                                                                       int max_tris, uint8_t* __restrict__ out, int H, int W,
                                                                       float bg_r, float bg_g, float bg_b) {
  __shared__ RastTri sm[RAST_CHUNK];
  const int e = blockIdx.z;
  const int px = blockIdx.x * RAST_TILE + threadIdx.x % RAST_TILE, py = blockIdx.y * RAST_TILE + threadIdx.x / RAST_TILE;
  const float tile_x0 = blockIdx.x * RAST_TILE, tile_y0 = blockIdx.y * RAST_TILE;
  const float cx = px + 0.5f, cy = py + 0.5f;
  float best_w = 0.f, cr = bg_r, cg = bg_g, cb = bg_b;  // larger w = nearer
  const int n = min(ntris[e], max_tris);
  const float* base = tris + (size_t)e * max_tris * 12;
  for (int c0 = 0; c0 < n; c0 += RAST_CHUNK) {
    const int cnt = min(RAST_CHUNK, n - c0);
    __syncthreads();
    for (int i = threadIdx.x; i < cnt * 12; i += blockDim.x) reinterpret_cast<float*>(sm)[i] = base[(size_t)c0 * 12 + i];
    __syncthreads();
    for (int t = 0; t < cnt; ++t) {
      const RastTri& T = sm[t];
      // tile-level bounding-box reject (uniform across the CTA → no divergence)
      const float minx = fminf(T.x0, fminf(T.x1, T.x2)), maxx = fmaxf(T.x0, fmaxf(T.x1, T.x2));
      const float miny = fminf(T.y0, fminf(T.y1, T.y2)), maxy = fmaxf(T.y0, fmaxf(T.y1, T.y2));
      if (maxx < tile_x0 || minx > tile_x0 + RAST_TILE || maxy < tile_y0 || miny > tile_y0 + RAST_TILE) continue;
      const float area = (T.x1 - T.x0) * (T.y2 - T.y0) - (T.x2 - T.x0) * (T.y1 - T.y0);
      if (fabsf(area) < 1e-9f) continue;
      const float e0 = (T.x1 - cx) * (T.y2 - cy) - (T.x2 - cx) * (T.y1 - cy);
      const float e1 = (T.x2 - cx) * (T.y0 - cy) - (T.x0 - cx) * (T.y2 - cy);
      const float e2 = (T.x0 - cx) * (T.y1 - cy) - (T.x1 - cx) * (T.y0 - cy);
      const float inv = 1.f / area;
      const float b0 = e0 * inv, b1 = e1 * inv, b2 = e2 * inv;
      if (b0 < 0.f || b1 < 0.f || b2 < 0.f) continue;        // either winding: barycentrics are sign-normalised by area
      const float w = b0 * T.w0 + b1 * T.w1 + b2 * T.w2;      // 1/z interpolates linearly in screen space
      if (w > best_w) {
        best_w = w;
        cr = T.r; cg = T.g; cb = T.b;
      }
    }
  }
  if (px < W && py < H) {
    uint8_t* o = out + (((size_t)e * H + py) * W + px) * 3;
    o[0] = (uint8_t)fminf(fmaxf(cr, 0.f), 255.f);
    o[1] = (uint8_t)fminf(fmaxf(cg, 0.f), 255.f);
    o[2] = (uint8_t)fminf(fmaxf(cb, 0.f), 255.f);
  }
}

// tris: fp32 [E, max_tris, 12] (x0 y0 w0 x1 y1 w1 x2 y2 w2 r g b), ntris: int32 [E]; out: uint8 [E, H, W, 3]
RT1_API int rt1_raster(const float* tris, const int* ntris, int max_tris, unsigned char* out, int E, int H, int W, float bg_r,
                       float bg_g, float bg_b, cudaStream_t stream) {
  dim3 grid(rt1_cdiv(W, RAST_TILE), rt1_cdiv(H, RAST_TILE), E);
  raster_kernel<<<grid, RAST_TILE * RAST_TILE, 0, stream>>>(tris, ntris, max_tris, out, H, W, bg_r, bg_g, bg_b);
  return rt1_check_launch("rt1_raster");
}

// Centre crop of `crop` (fraction of each side, e.g. 0.95) + bilinear resize (half-pixel centres, no antialias) of
// uint8 HWC frames to float32 CHW in [0, 1]: the policy's `image` observation.
__global__ void crop_resize_kernel(const uint8_t* __restrict__ in, float* __restrict__ out, int E, int H, int W, int OH, int OW,
                                   float y0, float x0, float sy, float sx) {
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const long total = (long)E * OH * OW;
  if (i >= total) return;
  const int ox = (int)(i % OW), oy = (int)((i / OW) % OH);
  const long e = i / ((long)OW * OH);
  const float fy = y0 + (oy + 0.5f) * sy - 0.5f, fx = x0 + (ox + 0.5f) * sx - 0.5f;
  const int iy0 = max(0, min(H - 1, (int)floorf(fy))), ix0 = max(0, min(W - 1, (int)floorf(fx)));
  const int iy1 = min(H - 1, iy0 + 1), ix1 = min(W - 1, ix0 + 1);
  const float ty = fminf(fmaxf(fy - iy0, 0.f), 1.f), tx = fminf(fmaxf(fx - ix0, 0.f), 1.f);
  const uint8_t* img = in + (size_t)e * H * W * 3;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const float v00 = img[((size_t)iy0 * W + ix0) * 3 + c], v01 = img[((size_t)iy0 * W + ix1) * 3 + c];
    const float v10 = img[((size_t)iy1 * W + ix0) * 3 + c], v11 = img[((size_t)iy1 * W + ix1) * 3 + c];
    const float v = (v00 * (1.f - tx) + v01 * tx) * (1.f - ty) + (v10 * (1.f - tx) + v11 * tx) * ty;
    out[(((size_t)e * 3 + c) * OH + oy) * OW + ox] = v * (1.f / 255.f);
  }
}

RT1_API int rt1_crop_resize(const unsigned char* in, float* out, int E, int H, int W, int OH, int OW, float crop, cudaStream_t stream) {
  const float ch = H * crop, cw = W * crop;
  const float y0 = (H - ch) * 0.5f, x0 = (W - cw) * 0.5f;
  const long total = (long)E * OH * OW;
  crop_resize_kernel<<<rt1_cdiv(total, 256), 256, 0, stream>>>(in, out, E, H, W, OH, OW, y0, x0, ch / OH, cw / OW);
  return rt1_check_launch("rt1_crop_resize");
}
