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

struct RastTri {
  float x0, y0, w0, x1, y1, w1, x2, y2, w2;  // screen-space vertices, w = 1 / camera depth (> 0 in front of the camera)
  float r, g, b;
};

__global__ void __launch_bounds__(RAST_TILE * RAST_TILE) raster_kernel(const float* __restrict__ tris, const int* __restrict__ ntris,
                                                                       int max_tris, uint8_t* __restrict__ out, int H, int W,
                                                                       float bg_r, float bg_g, float bg_b) {
  // Per chunk of 256 triangles: thread t tests triangle t against this CTA's 16x16 tile (bounding box, degenerate area) and
  // the survivors are compacted into shared memory (warp ballot + one shared counter); only they enter the per-pixel loop.
  // With the reference's meshes a board is ~2 000 triangles of which a tile touches a few dozen — without the compaction
  // every thread walked all of them (5 ms per 64 frames instead of 0.6 ms with the procedural prisms).
  __shared__ RastTri sm[RAST_TILE * RAST_TILE];
  __shared__ int n_keep;
  const int e = blockIdx.z;
  const int px = blockIdx.x * RAST_TILE + threadIdx.x % RAST_TILE, py = blockIdx.y * RAST_TILE + threadIdx.x / RAST_TILE;
  const float tile_x0 = blockIdx.x * RAST_TILE, tile_y0 = blockIdx.y * RAST_TILE;
  const float cx = px + 0.5f, cy = py + 0.5f;
  float best_w = 0.f, cr = bg_r, cg = bg_g, cb = bg_b;  // larger w = nearer
  const int n = min(ntris[e], max_tris);
  const float* base = tris + (size_t)e * max_tris * 12;
  const int lane = threadIdx.x & 31;
  for (int c0 = 0; c0 < n; c0 += RAST_TILE * RAST_TILE) {
    __syncthreads();
    if (threadIdx.x == 0) n_keep = 0;
    __syncthreads();
    const int ti = c0 + threadIdx.x;
    RastTri T;
    bool keep = false;
    if (ti < n) {
      const float4* p4 = reinterpret_cast<const float4*>(base + (size_t)ti * 12);   // 48-byte records: 16-byte aligned
      const float4 a = p4[0], b = p4[1], c = p4[2];
      T.x0 = a.x; T.y0 = a.y; T.w0 = a.z; T.x1 = a.w; T.y1 = b.x; T.w1 = b.y; T.x2 = b.z; T.y2 = b.w; T.w2 = c.x;
      T.r = c.y; T.g = c.z; T.b = c.w;
      const float minx = fminf(T.x0, fminf(T.x1, T.x2)), maxx = fmaxf(T.x0, fmaxf(T.x1, T.x2));
      const float miny = fminf(T.y0, fminf(T.y1, T.y2)), maxy = fmaxf(T.y0, fmaxf(T.y1, T.y2));
      const float area = (T.x1 - T.x0) * (T.y2 - T.y0) - (T.x2 - T.x0) * (T.y1 - T.y0);
      keep = !(maxx < tile_x0 || minx > tile_x0 + RAST_TILE || maxy < tile_y0 || miny > tile_y0 + RAST_TILE) && fabsf(area) >= 1e-9f;
    }
    // order-preserving within a warp; across warps the order is irrelevant (nearest hit wins, ties are measure-zero)
    const unsigned m = __ballot_sync(0xffffffffu, keep);
    int warp_base = 0;
    if (lane == 0 && m) warp_base = atomicAdd(&n_keep, __popc(m));
    warp_base = __shfl_sync(0xffffffffu, warp_base, 0);
    if (keep) sm[warp_base + __popc(m & ((1u << lane) - 1u))] = T;
    __syncthreads();
    const int cnt = n_keep;
    for (int t = 0; t < cnt; ++t) {
      const RastTri& Q = sm[t];
      const float area = (Q.x1 - Q.x0) * (Q.y2 - Q.y0) - (Q.x2 - Q.x0) * (Q.y1 - Q.y0);
      const float e0 = (Q.x1 - cx) * (Q.y2 - cy) - (Q.x2 - cx) * (Q.y1 - cy);
      const float e1 = (Q.x2 - cx) * (Q.y0 - cy) - (Q.x0 - cx) * (Q.y2 - cy);
      const float e2 = (Q.x0 - cx) * (Q.y1 - cy) - (Q.x1 - cx) * (Q.y0 - cy);
      const float inv = 1.f / area;
      const float b0 = e0 * inv, b1 = e1 * inv, b2 = e2 * inv;
      if (b0 < 0.f || b1 < 0.f || b2 < 0.f) continue;        // either winding: barycentrics are sign-normalised by area
      const float w = b0 * Q.w0 + b1 * Q.w1 + b2 * Q.w2;      // 1/z interpolates linearly in screen space
      if (w > best_w) {
        best_w = w;
        cr = Q.r; cg = Q.g; cb = Q.b;
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
  RT1_REQUIRE((reinterpret_cast<uintptr_t>(tris) & 15) == 0, "rt1_raster: tris must be 16-byte aligned (float4 record loads)");
  dim3 grid(rt1_cdiv(W, RAST_TILE), rt1_cdiv(H, RAST_TILE), E);
  raster_kernel<<<grid, RAST_TILE * RAST_TILE, 0, stream>>>(tris, ntris, max_tris, out, H, W, bg_r, bg_g, bg_b);
  return rt1_check_launch("rt1_raster");
}

// Centre crop of `crop` (fraction of each side, e.g. 0.95; whole pixels, see rt1_crop_resize) + bilinear resize (half-pixel centres, no antialias) of
// uint8 HWC frames to float32 CHW in [0, 1]: the policy's `image` observation.
__global__ void crop_resize_kernel(const uint8_t* __restrict__ in, float* __restrict__ out, int E, int H, int W, int OH, int OW,
                                   float y0, float x0, float sy, float sx, int ylo, int yhi, int xlo, int xhi) {
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const long total = (long)E * OH * OW;
  if (i >= total) return;
  const int ox = (int)(i % OW), oy = (int)((i / OW) % OH);
  const long e = i / ((long)OW * OH);
  const float fy = y0 + (oy + 0.5f) * sy - 0.5f, fx = x0 + (ox + 0.5f) * sx - 0.5f;
  // the crop window is its own image: samples are clamped to it, like a resize of the cropped frame
  const int iy0 = max(ylo, min(yhi, (int)floorf(fy))), ix0 = max(xlo, min(xhi, (int)floorf(fx)));
  const int iy1 = min(yhi, iy0 + 1), ix1 = min(xhi, ix0 + 1);
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
  // the reference's evaluation crop (language_table/eval/wrappers.py:99-118): a window of int(H f) x int(W f) pixels at the
  // integer offset ((H - H f) // 2, (W - W f) // 2), then a resize of that window
  const float ch = floorf(H * crop), cw = floorf(W * crop);
  const float y0 = floorf((H - H * crop) * 0.5f), x0 = floorf((W - W * crop) * 0.5f);
  const long total = (long)E * OH * OW;
  const int ylo = (int)y0, xlo = (int)x0, yhi = ylo + (int)ch - 1, xhi = xlo + (int)cw - 1;
  crop_resize_kernel<<<rt1_cdiv(total, 256), 256, 0, stream>>>(in, out, E, H, W, OH, OW, y0, x0, ch / OH, cw / OW, ylo, yhi < H ? yhi : H - 1,
                                                               xlo, xhi < W ? xhi : W - 1);
  return rt1_check_launch("rt1_crop_resize");
}

// Per-frame crop box + bilinear resize of raw uint8 HWC frames — the GPU half of the training input pipeline.  The reference
// does a PIL crop of 95 % of the frame at a random offset and a bilinear resize to 456 x 256 per frame on the CPU workers
// (load_np_dataset.py:8-39); here the workers ship the raw frames plus one box per frame and this kernel produces the
// network's [E, 3, OH, OW] input.  Sampling: half-pixel centres, no antialiasing (identical to PIL when enlarging, which is
// the Language-Table case: 180 x 320 → 256 x 456).  boxes fp32 [E, 4] = (top, left, height, width) in source pixels.
template <typename OUT>
__global__ void crop_resize_boxes_kernel(const uint8_t* __restrict__ in, const float* __restrict__ boxes, OUT* __restrict__ out,
                                         int E, int H, int W, int OH, int OW) {
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const long total = (long)E * OH * OW;
  if (i >= total) return;
  const int ox = (int)(i % OW), oy = (int)((i / OW) % OH);
  const long e = i / ((long)OW * OH);
  const float4 box = reinterpret_cast<const float4*>(boxes)[e];
  const float sy = box.z / OH, sx = box.w / OW;
  const float fy = box.x + (oy + 0.5f) * sy - 0.5f, fx = box.y + (ox + 0.5f) * sx - 0.5f;
  // the box is its own image: samples are clamped to it, like a resize of the cropped frame
  const int ylo = max(0, (int)floorf(box.x)), yhi = min(H - 1, (int)ceilf(box.x + box.z) - 1);
  const int xlo = max(0, (int)floorf(box.y)), xhi = min(W - 1, (int)ceilf(box.y + box.w) - 1);
  const int iy0 = max(ylo, min(yhi, (int)floorf(fy))), ix0 = max(xlo, min(xhi, (int)floorf(fx)));
  const int iy1 = min(yhi, iy0 + 1), ix1 = min(xhi, ix0 + 1);
  const float ty = fminf(fmaxf(fy - iy0, 0.f), 1.f), tx = fminf(fmaxf(fx - ix0, 0.f), 1.f);
  const uint8_t* img = in + (size_t)e * H * W * 3;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const float v00 = img[((size_t)iy0 * W + ix0) * 3 + c], v01 = img[((size_t)iy0 * W + ix1) * 3 + c];
    const float v10 = img[((size_t)iy1 * W + ix0) * 3 + c], v11 = img[((size_t)iy1 * W + ix1) * 3 + c];
    const float v = (v00 * (1.f - tx) + v01 * tx) * (1.f - ty) + (v10 * (1.f - tx) + v11 * tx) * ty;
    const size_t o = (((size_t)e * 3 + c) * OH + oy) * OW + ox;
    if (sizeof(OUT) == 1) out[o] = (OUT)(v + 0.5f);
    else out[o] = (OUT)(v * (1.f / 255.f));
  }
}

// in uint8 [E, H, W, 3]; boxes fp32 [E, 4] (16-byte aligned); out uint8 (out_uint8 = 1) or fp32 in [0, 1], [E, 3, OH, OW]
RT1_API int rt1_crop_resize_boxes(const unsigned char* in, const float* boxes, void* out, int out_uint8, int E, int H, int W, int OH,
                                  int OW, cudaStream_t stream) {
  RT1_REQUIRE(((uintptr_t)boxes & 15) == 0, "rt1_crop_resize_boxes: boxes must be 16-byte aligned");
  const long total = (long)E * OH * OW;
  if (total <= 0) return 0;
  if (out_uint8) crop_resize_boxes_kernel<uint8_t><<<rt1_cdiv(total, 256), 256, 0, stream>>>(in, boxes, (uint8_t*)out, E, H, W, OH, OW);
  else crop_resize_boxes_kernel<float><<<rt1_cdiv(total, 256), 256, 0, stream>>>(in, boxes, (float*)out, E, H, W, OH, OW);
  return rt1_check_launch("rt1_crop_resize_boxes");
}
